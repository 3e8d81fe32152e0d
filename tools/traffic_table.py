"""Per-kernel DRAM traffic of one fold from an ncu CSV with dram__bytes_read.sum / dram__bytes_write.sum
per launch (tools/evidence.sh, <tag>_dram.csv).  usage: traffic_table.py <dram.csv> [fold-index]"""
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
H = rows[0]
ki, mi, ui, vi, ii = (H.index(x) for x in ('Kernel Name', 'Metric Name', 'Metric Unit', 'Metric Value', 'ID'))
scale = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
launch = {}
for r in rows[1:]:
    d = launch.setdefault(int(r[ii]), [r[ki].split('(')[0].replace('void ', ''), 0.0, 0.0])
    v = float(r[vi].replace(',', '')) * scale.get(r[ui], 1.0)
    if 'read' in r[mi]: d[1] += v
    else: d[2] += v
seq = [launch[k] for k in sorted(launch)]
inits = [i for i, s in enumerate(seq) if s[0] == 'k_init'] + [len(seq)]
f = int(sys.argv[2]) if len(sys.argv) > 2 else 0
tot = {}
for k, rd, wr in seq[inits[f]:inits[f + 1]]:
    t = tot.setdefault(k, [0, 0.0, 0.0]); t[0] += 1; t[1] += rd; t[2] += wr
print("%-30s %8s %12s %12s %16s" % ("kernel", "launches", "read MB", "write MB", "MB per launch"))
for k, (n, rd, wr) in sorted(tot.items(), key=lambda kv: -(kv[1][1] + kv[1][2])):
    print("%-30s %8d %12.1f %12.1f %16.2f" % (k[:30], n, rd / 1e6, wr / 1e6, (rd + wr) / n / 1e6))
print("%-30s %8s %12.1f %12.1f" % ("fold total", "", sum(t[1] for t in tot.values()) / 1e6, sum(t[2] for t in tot.values()) / 1e6))
