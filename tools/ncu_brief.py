"""Brief per-kernel table of an ncu --set full report.  usage: ncu_brief.py <report.ncu-rep>"""
import csv, io, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); H = rows[0]; U = rows[1]
pats = ['Kernel Name', 'gpu__time_duration.sum', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'l1tex__t_sector_hit_rate.pct',
        'lts__t_sector_hit_rate.pct', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active']
for i, h in enumerate(H):
    if h in pats or (h.startswith('smsp__average_warps_issue_stalled') and 'per_issue_active' in h):
        vals = [r[i] for r in rows[2:]]
        try:
            if all(float(v.replace(',', '')) < 0.1 for v in vals): continue
        except Exception: pass
        print("%-80s %-8s %s" % (h[:80], U[i][:8], ' | '.join(v[:16] for v in vals)))
