"""Key figures of one `ncu --page raw --csv` export: python tools/ncu_key.py file.csv"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1], errors="replace")))
hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
H, U = rows[hi], rows[hi + 1]
for V in rows[hi + 2:]:
    d = dict(zip(H, V))
    u = dict(zip(H, U))
    def g(k):
        return d.get(k, "?") + " " + u.get(k, "")
    print(d["Kernel Name"][:90], "grid", d.get("Grid Size"), "block", d.get("Block Size"))
    for k in ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes.sum.per_second", "launch__registers_per_thread",
              "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
              "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "smsp__average_warp_latency_per_inst_issued.ratio", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
              "smsp__warps_active.avg.per_cycle_active", "launch__waves_per_multiprocessor"]:
        if k in d: print("   %-70s %s" % (k, g(k)))
    st = [(float(d[k].replace(",", "")), k) for k in d if "issue_stalled" in k and k.endswith("per_issue_active.ratio") and d[k] not in ("", "n/a")]
    for v, k in sorted(st, reverse=True)[:6]:
        print("   stall %-40s %.2f" % (k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v))
