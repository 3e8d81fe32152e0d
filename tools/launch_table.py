"""Per-folding-iteration kernel times of one fold from an ncu launch list (gpu__time_duration).
usage: launch_table.py <launches.csv> [fold-index]"""
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
H = rows[0]; ki = H.index('Kernel Name'); vi = H.index('Metric Value')
seq = [(r[ki].split('(')[0].replace('void ', ''), float(r[vi].replace(',', ''))) for r in rows[1:]]
inits = [i for i, s in enumerate(seq) if s[0] == 'k_init'] + [len(seq)]
f = int(sys.argv[2]) if len(sys.argv) > 2 else 0
s = seq[inits[f]:inits[f + 1]]
it, line, tot = 0, [], {}
for k, v in s:
    if k == 'k_gather':
        if line: print(it, ' '.join(line)); it += 1
        line = []
    if v > 5000: line.append("%s=%.0f" % (k.replace('k_', '')[:14], v / 1e3))
    tot[k] = tot.get(k, 0) + v
print(it, ' '.join(line))
T = sum(tot.values())
print("total %.2f ms" % (T / 1e6))
for k, v in sorted(tot.items(), key=lambda kv: -kv[1]): print("  %-28s %8.3f ms %5.1f%%" % (k[:28], v / 1e6, 100 * v / T))
