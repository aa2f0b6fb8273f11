"""Summarises an .ncu-rep: key raw metrics per launch and the hottest source lines (needs -lineinfo + --import-source)."""
import csv, subprocess, sys, io, collections
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active","sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio"]
for k in want:
    if k in hdr:
        i = hdr.index(k)
        print(f"{k:80s} {units[i]:12s} " + " | ".join(r[i][:40] for r in rows[2:]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"] , capture_output=True, text=True).stdout
# split per kernel launch: take the last one
blocks = src.split('"Address"')
if len(blocks) > 1:
    last = '"Address"' + blocks[-1]
    rows = list(csv.reader(io.StringIO(last)))
    h = rows[0]
    def col(name):
        for i, c in enumerate(h):
            if c == name: return i
        return None
    ci, cs, cx = col("Source"), col("Warp Stall Sampling (All Samples)"), col("# Instructions Executed") or col("Instructions Executed")
    print("columns:", h[:12])
    tot = 0; data = []
    for r in rows[1:]:
        try: s = int(r[cs]); 
        except Exception: continue
        tot += s; data.append((s, r))
    print("total samples", tot)
    for s, r in sorted(data, key=lambda x: -x[0])[:top]:
        print(f"{s:7d} {100*s/max(tot,1):5.1f}%  {r[ci][:110]}")
