"""Raw-page extract of an ncu report: python scratch/ncu_extract.py <report.ncu-rep> <out.csv> -- one row per captured launch
with the metrics profiles/README.md quotes (time, DRAM bytes, L1/L2 hit rates, issue rate, occupancy, global-load sectors per
request, shared-memory bank conflicts, stall ratios)."""
import csv, subprocess, sys
rep, out = sys.argv[1], sys.argv[2]
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ["Kernel Name", "Grid Size", "Block Size", "launch__cluster_dim_x", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__cycles_active.avg", "sm__cycles_active.sum", "launch__registers_per_thread", "launch__occupancy_limit_shared_mem",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "smsp__inst_executed_op_shared_atom.sum",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio", "sm__ctas_launched.sum"]
cols = [(w, hdr.index(w)) for w in want if w in hdr]
num = lambda v: float(v.replace(",", ""))
with open(out, "w", newline="") as f:
    w = csv.writer(f)
    w.writerow([c if not units[i] else f"{c} [{units[i]}]" for c, i in cols] + ["global_load_sectors_per_request"])
    for r in data:
        d = {c: r[i] for c, i in cols}
        try:
            spr = num(d["l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum"]) / num(d["l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum"])
        except Exception:
            spr = ""
        d["Kernel Name"] = d["Kernel Name"].split("(")[0]
        w.writerow([d[c] for c, _ in cols] + [spr])
print(len(data), "launches ->", out)
