"""Summarise an `ncu --page raw --csv` dump: python profiles/ncu_summary.py file.raw.csv [regex]"""
import csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
pat = re.compile(sys.argv[2] if len(sys.argv) > 2 else
                 r"^(gpu__time_duration.sum|dram__bytes_(read|write).sum|lts__t_bytes.sum|lts__t_sector_hit_rate.pct|"
                 r"l1tex__t_sector_hit_rate.pct|smsp__inst_executed.sum|sm__warps_active.avg.pct_of_peak_sustained_active|"
                 r"smsp__issue_active.avg.pct_of_peak_sustained_active|sm__throughput.avg.pct_of_peak_sustained_elapsed|"
                 r"gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed|launch__registers_per_thread|launch__grid_size|"
                 r"launch__block_size|launch__occupancy_limit_.*|smsp__warp_issue_stalled_.*_per_warp_active.pct|"
                 r"smsp__average_warp_latency_per_inst_issued.ratio|smsp__thread_inst_executed_per_inst_executed.ratio|"
                 r"smsp__inst_executed_pipe_.*\.sum|sm__inst_executed_pipe_.*\.sum|smsp__warps_eligible.avg.per_cycle_active|"
                 r"l1tex__data_pipe_lsu_wavefronts.sum|lts__t_sectors_srcunit_tex_op_read.sum|sm__cycles_elapsed.max|"
                 r"l1tex__t_bytes_pipe_lsu_mem_global_op_ld.sum|smsp__inst_executed_op_.*)$")
for d in data:
    name = d[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
    print("== launch", d[0], name)
    for i, k in enumerate(hdr):
        if pat.match(k) and d[i] not in ("", "0", "n/a"):
            print(f"  {k:75s} {d[i]:>18s} {units[i]}")
