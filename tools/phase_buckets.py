"""Bucket the NCU_LINES_INNER=1 output of ncu_lines.py by expansion phase (function line ranges are
read from expand_fast.cuh).  usage: NCU_LINES_INNER=1 ncu_lines.py ... 4000 | phase_buckets.py"""
import re, sys, os
src = open(os.path.join(os.path.dirname(__file__), "..", "rafft_b200", "csrc", "expand_fast.cuh")).read().splitlines()
marks = []
for i, l in enumerate(src, 1):
    m = re.match(r"__device__ \S+ (?:\w+ )?(\w+)\(", l) or re.match(r"__device__ __forceinline__ (?:unsigned|int|void) (\w+)\(", l)
    if m: marks.append((i, m.group(1)))
fc = open(os.path.join(os.path.dirname(__file__), "..", "rafft_b200", "csrc", "fold_core.cuh")).read().splitlines()
def _at(pat):
    return next(i for i, l in enumerate(fc, 1) if pat in l)
_m = [(_at("RF_FN uint64_t mix64"), "pair_fp/mix64"), (_at("RF_FN int pair_type"), "pair_type"), (_at("RF_FN void cand_geometry(Cand &c, const PosT *pos, const BrT *brcnt, int mip, int mjp, int nb) {"), "cand_geometry"),
      (_at("struct EnergyView"), "energy view"), (_at("RF_NOINLINE int e_hairpin"), "e_hairpin"), (_at("RF_NOINLINE int e_interior"), "e_interior"),
      (_at("RF_FN int e_mlstem"), "ml/ext stem"), (_at("RF_NOINLINE int loop_energy"), "loop_energy"), (_at("struct alignas(16) SeenEnt"), None)]
FC = [(a, _m[k + 1][0], nm) for k, (a, nm) in enumerate(_m[:-1]) if nm]
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
        key = next((nm for a, b, nm in FC if a <= ln < b), 'fold_core %d' % ln)
    elif f == 'rafft_cuda.cu': key = 'kernel body (barriers, fetch)'
    a = agg.setdefault(key, [0, 0, 0]); a[0] += i; a[1] += s; a[2] += i * th
for k, (i, s, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]): print('%-32s inst %6.2f%%  samples %6.2f%%  active threads %4.1f' % (k, i, s, t / max(i, 1e-9)))
