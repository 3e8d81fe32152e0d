"""Static SASS size of a kernel (and its callees) attributed to the INNERMOST repo source line.
usage: sass_lines.py <lib.so> <kernel-substring> [bucket]   (bucket = group lines into ranges of N)"""
import collections, os, re, subprocess, sys, tempfile
lib, kern = sys.argv[1:3]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
insec, group, fresh, func = False, [], True, "kernel"
byline = collections.Counter()
for line in dis.splitlines():
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", line)
    if m:
        insec = kern in m.group(1); continue
    if not insec: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        if fresh: group = []; fresh = False
        group.append((os.path.basename(m.group(1)), int(m.group(2)))); continue
    if re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+\S", line):
        fresh = True
        pick = next((g for g in group if g[0] in ("expand_fast.cuh", "fold_core.cuh", "rafft_cuda.cu")), group[0] if group else ("?", 0))
        byline[pick] += 1
print("total", sum(byline.values()))
for k, c in sorted(byline.items()):
    if c >= 8: print("%6d  %s:%s" % (c, k[0], k[1]))
