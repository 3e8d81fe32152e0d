"""Bucket the NCU_LINES_INNER=1 output of ncu_lines.py by expansion phase (function line ranges are
read from expand_fast.cuh).  usage: NCU_LINES_INNER=1 ncu_lines.py ... 4000 | phase_buckets.py"""
import re, sys, os
src = open(os.path.join(os.path.dirname(__file__), "..", "rafft_b200", "csrc", "expand_fast.cuh")).read().splitlines()
marks = []
for i, l in enumerate(src, 1):
    m = re.match(r"__device__ \S+ (?:\w+ )?(\w+)\(", l) or re.match(r"__device__ __forceinline__ (?:unsigned|int|void) (\w+)\(", l)
    if m: marks.append((i, m.group(1)))
def fn(ln):
    name = "?"
    for a, n in marks:
        if a <= ln + 1: name = n
    return name
agg = {}
for l in sys.stdin:
    m = re.match(r'(\S+):(\d+)\s+inst\s+([\d.]+)%\s+samples\s+([\d.]+)%\s+thr\s+([\d.]+)', l)
    if not m: continue
    f, ln, i, s, th = m.group(1), int(m.group(2)), float(m.group(3)), float(m.group(4)), float(m.group(5))
    key = f
    if f == 'expand_fast.cuh': key = fn(ln)
    elif f == 'fold_core.cuh':
        key = "pair_fp/mix64" if 297 <= ln <= 307 else "pair_type" if 308 <= ln <= 314 else "e_hairpin" if 323 <= ln <= 352 else "e_interior" if 354 <= ln <= 390 else "ml/ext stem" if 392 <= ln <= 406 else 'fold_core %d' % ln
    elif f == 'rafft_cuda.cu': key = 'kernel body (barriers, fetch)'
    a = agg.setdefault(key, [0, 0, 0]); a[0] += i; a[1] += s; a[2] += i * th
for k, (i, s, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]): print('%-32s inst %6.2f%%  samples %6.2f%%  active threads %4.1f' % (k, i, s, t / max(i, 1e-9)))
