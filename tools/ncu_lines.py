"""Attribute an ncu source-page CSV (SASS rows) to CUDA source lines using nvdisasm line info.

usage: ncu_lines.py <report.ncu-rep> <lib.so> <kernel-substring> [top]
Prints per-source-line totals of executed warp instructions and stall samples."""
import collections, csv, io, os, re, subprocess, sys, tempfile

rep, lib, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
if rep.endswith(".gz"):  # source page already exported on the GPU box (tools/evidence.sh)
    import gzip
    src = gzip.open(rep, "rt").read()
else:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
H = rows[hi]
data = [dict(zip(H, r)) for r in rows[hi + 1:] if len(r) == len(H)]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
insts, cur, insec = [], None, False
group, fresh = [], True
INNER = os.environ.get('NCU_LINES_INNER', '0') == '1'
for line in dis.splitlines():
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", line)
    if m:
        insec = kern in m.group(1)
        continue
    if not insec:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        if fresh:
            group = []; fresh = False
        group.append((os.path.basename(m.group(1)), int(m.group(2))))
        cur = group[-1]
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m:
        fresh = True
        pick = cur
        if INNER:  # innermost line that lies in this repository's sources
            pick = next((g for g in group if not g[0].endswith((".hpp", ".h"))), cur)
        insts.append((pick, m.group(2).strip()))
print("sass rows: ncu %d, nvdisasm %d" % (len(data), len(insts)))
n = min(len(data), len(insts))
mism = sum(1 for k in range(n) if data[k]["Source"].split()[0].lstrip("@!P0123456789T ") [:3] != insts[k][1].split()[0].lstrip("@!P0123456789T ")[:3])
print("opcode mismatches:", mism)
agg = collections.defaultdict(lambda: [0, 0, collections.Counter(), 0])
stall_cols = [h for h in H if h.startswith("stall_")]
for k in range(n):
    d = data[k]
    a = agg[insts[k][0]]
    a[0] += int(d["Instructions Executed"] or 0)
    a[1] += int(d["# Samples"] or 0)
    a[3] += int(d.get("Predicated-On Thread Instructions Executed") or 0)
    for s in stall_cols:
        v = int(d[s] or 0)
        if v:
            a[2][s[6:]] += v
ti = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values())
print("total warp inst %d, samples %d" % (ti, ts))
print("--- by samples")
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print("%-22s inst %6.2f%%  samples %6.2f%%  thr %5.1f  %s" % ("%s:%d" % key if key else "?", 100.0 * a[0] / ti, 100.0 * a[1] / ts, a[3] / max(a[0], 1),
                                                  " ".join("%s=%d" % kv for kv in a[2].most_common(3))))
