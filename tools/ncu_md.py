"""Markdown table of the metrics that matter from `ncu --page raw --csv` exports (or .ncu-rep files),
one column per report.  usage: ncu_md.py name=file [name=file ...]"""
import csv, io, subprocess, sys
M = [("time", "gpu__time_duration.sum"), ("warp instructions", "smsp__inst_executed.sum"),
     ("issue slots busy %", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
     ("active threads / instruction", "smsp__thread_inst_executed_per_inst_executed.ratio"),
     ("achieved occupancy %", "sm__warps_active.avg.pct_of_peak_sustained_active"),
     ("ALU pipe %", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
     ("FP64 pipe %", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"),
     ("DRAM read", "dram__bytes_read.sum"), ("DRAM write", "dram__bytes_write.sum"),
     ("L1 hit %", "l1tex__t_sector_hit_rate.pct"), ("L2 hit %", "lts__t_sector_hit_rate.pct"),
     ("shared-memory bank conflicts", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"),
     ("registers", "launch__registers_per_thread"), ("block size", "launch__block_size"), ("grid size", "launch__grid_size"),
     ("dynamic shared memory / block", "launch__shared_mem_per_block_dynamic")] + \
    [("stall %s / issue" % s, "smsp__average_warps_issue_stalled_%s_per_issue_active.ratio" % s)
     for s in ("long_scoreboard", "short_scoreboard", "barrier", "no_instruction", "wait", "math_pipe_throttle",
               "branch_resolving", "not_selected", "lg_throttle", "mio_throttle")]
cols = []
for a in sys.argv[1:]:
    name, f = a.split("=", 1)
    raw = open(f).read() if f.endswith(".csv") else subprocess.run(["ncu", "-i", f, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = [r for r in csv.reader(io.StringIO(raw)) if len(r) > 20]
    H, U, V = rows[0], rows[1], rows[2]
    cols.append((name, {h: (U[i], V[i]) for i, h in enumerate(H)}))
def fmt(v):
    try: return "%.3g" % float(v.replace(",", ""))
    except ValueError: return v
print("| metric | " + " | ".join(n for n, _ in cols) + " |")
print("|---|" + "---|" * len(cols))
for label, key in M:
    if not any(key in d for _, d in cols): continue
    cells = ["%s %s" % (fmt(d[key][1]), d[key][0]) if key in d else "-" for _, d in cols]
    print("| %s (`%s`) | " % (label, key) + " | ".join(c.strip() for c in cells) + " |")
