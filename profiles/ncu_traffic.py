"""python profiles/ncu_traffic.py gpurun_out/k1_TAG.raw.csv TAG  ->  profiles/k1_traffic.json
DRAM bytes (read + write) of the captured K1 launch, for bench.py's roofline.traffic."""
import csv, json, os, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, d = rows[0], rows[1], rows[2]
def val(name):
    i = hdr.index(name)
    v, u = float(d[i].replace(",", "")), units[i]
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
out = {"dram_bytes_per_launch": val("dram__bytes_read.sum") + val("dram__bytes_write.sum"),
       "dram_bytes_read": val("dram__bytes_read.sum"), "dram_bytes_write": val("dram__bytes_write.sum"),
       "kernel": d[hdr.index("Kernel Name")], "gpu_time_us": float(d[hdr.index("gpu__time_duration.sum")].replace(",", "")),
       "source": f"ncu --set full capture {sys.argv[2]} of bench.py's K1 launch (cold L2: 256 MiB flushed before it)"}
json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "k1_traffic.json"), "w"), indent=1)
print(out)
