"""Summarise an ncu --set full report (raw page CSV) into the few metrics the roofline discussion uses.
usage: ncu -i rep --page raw --csv > raw.csv; python tools/ncu_summary.py raw.csv out.json"""
import csv, json, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
keep = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__cycles_elapsed.max", "sm__cycles_elapsed.avg.per_second", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__occupancy_limit_shared_mem", "lts__t_sector_hit_rate.pct",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__m_xbar2l1tex_read_bytes.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sass__inst_executed_local_loads", "sass__inst_executed_local_stores"]
out = []
ki = hdr.index("Kernel Name")
for r in rows[2:]:
    d = {"kernel": r[ki]}
    for h, u, v in zip(hdr, units, r):
        if h in keep:
            d[h + (f" [{u}]" if u else "")] = v
    out.append(d)
json.dump(out, open(sys.argv[2], "w"), indent=1)
for d in out:
    print(d["kernel"][:70], {k.split(".")[0][:28]: v for k, v in d.items() if k != "kernel" and ("duration" in k or "dram__bytes" in k or "tensor" in k)})
