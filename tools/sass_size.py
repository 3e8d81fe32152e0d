"""Static SASS size of a kernel attributed to (outermost) source lines / functions."""
import collections, os, re, subprocess, sys, tempfile
lib, kern = sys.argv[1:3]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
insec, cur, func = False, None, None
byline, byfunc = collections.Counter(), collections.Counter()
for line in dis.splitlines():
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", line)
    if m:
        insec = kern in m.group(1); func = "kernel"; continue
    if not insec: continue
    m = re.match(r"^(\$?[\w$]+):\s*$", line)
    if m and not m.group(1).startswith(".L"):
        func = m.group(1); 
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+\S", line):
        byline[cur] += 1; byfunc[func] += 1
print("total", sum(byline.values()))
for f, c in byfunc.most_common(20): print("%6d  %s" % (c, f[:90]))
print("--- lines")
for k, c in byline.most_common(40): print("%6d  %s:%s" % (c, k[0], k[1]))
