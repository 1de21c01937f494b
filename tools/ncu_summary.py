"""Markdown table of the metrics DESIGN.md quotes from an ncu report: python tools/ncu_summary.py <report.ncu-rep>"""
import csv, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h, units = rows[0], rows[1]
cols = [("ms", "gpu__time_duration.sum"), ("DMMA pipe %", "sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active"),
        ("FP64 pipe %", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
        ("SM thr %", "sm__throughput.avg.pct_of_peak_sustained_elapsed"), ("issue %", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
        ("DRAM rd", "dram__bytes_read.sum"), ("DRAM wr", "dram__bytes_write.sum"), ("L2 hit %", "lts__t_sector_hit_rate.pct"),
        ("regs", "launch__registers_per_thread"), ("grid", "launch__grid_size"), ("warps/SMSP", "smsp__warps_active.avg.per_cycle_active"),
        ("lanes/inst", "smsp__thread_inst_executed_per_inst_executed.ratio"),
        ("st.noinst", "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio"),
        ("st.wait", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio"),
        ("st.math", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio"),
        ("st.barrier", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio"),
        ("st.long_sb", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio")]
print("| kernel | " + " | ".join(c for c, _ in cols) + " |")
print("|---|" + "---|" * len(cols))
for r in rows[2:]:
    d = dict(zip(h, r)); u = dict(zip(h, units))
    cells = []
    for c, m in cols:
        v = d.get(m, "")
        try:
            f = float(v.replace(",", ""))
            if m == "gpu__time_duration.sum": f = f / 1e6 if u[m] in ("ns", "nsecond") else (f / 1e3 if u[m].startswith("u") else f)
            v = ("%.2f" % f) if abs(f) < 1000 else ("%.3g" % f)
            if m.startswith("dram__bytes"): v += " " + u[m]
        except ValueError:
            pass
        cells.append(v)
    print("| " + d["Kernel Name"].split("(")[0] + " | " + " | ".join(cells) + " |")
