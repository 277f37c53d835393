"""python profiles/ncu_traffic.py gpurun_out/k1_TAG.raw.csv TAG [out.json]  ->  profiles/k1_traffic.json (or out.json)
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
       "kernel": d[hdr.index("Kernel Name")], "gpu_time_us": float(d[hdr.index("gpu__time_duration.sum")].replace(",", "")) *
       {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(units[hdr.index("gpu__time_duration.sum")], 1.0),
       "source": f"ncu --set full capture {sys.argv[2]} of one launch of that kernel inside bench.py (configs[2], one GPU)"}
name = sys.argv[3] if len(sys.argv) > 3 else "k1_traffic.json"
json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), name), "w"), indent=1)
print(out)
