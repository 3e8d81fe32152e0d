"""Per-sequence schedule of one fold of a benchmark batch (diagnostics): start/end device
timestamps -> makespan structure, duration vs length, concurrency over time."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
from rafft_b200 import rafft as R
from rafft_b200._native import COUNTER_NAMES as CN
from rafft_b200.synth import config_batch
wl = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
n, ms = (100, 50) if wl == "cfg2" else (200, 200)
seqs = config_batch("cfg2" if wl in ("cfg2", "cfg3") else wl, int(sys.argv[2]) if len(sys.argv) > 2 else None)
ctx = R.get_context(0)
offs = np.zeros(len(seqs) + 1, dtype=np.int64); np.cumsum([len(s) for s in seqs], out=offs[1:])
prm = R.make_params(n, ms, 1000, traj=False)
for _ in range(2):
    raw = ctx.fold_raw("".join(seqs).encode(), offs, prm)
c = raw.counters
t0, t1 = c[:, CN.index("t0_ns")], c[:, CN.index("t1_ns")]
L = np.diff(offs)
base = t0.min(); span = (t1.max() - base) / 1e6
dur = (t1 - t0) / 1e6
print("kernel_ms %.2f  makespan %.2f ms  sum(dur) %.1f ms  mean concurrency %.0f" % (raw.kernel_ms, span, dur.sum(), dur.sum() / span))
for lo, hi in ((0, 80), (80, 120), (120, 160), (160, 220), (220, 300), (300, 400), (400, 501)):
    m = (L >= lo) & (L < hi)
    if m.any():
        print("L[%3d,%3d) n=%5d  dur mean %.2f max %.2f ms  share of block-time %.3f  end max %.1f" %
              (lo, hi, m.sum(), dur[m].mean(), dur[m].max(), dur[m].sum() / dur.sum(), ((t1[m] - base) / 1e6).max()))
edges = np.linspace(0, span, 21)
conc = [(((t0 - base) / 1e6 < b) & ((t1 - base) / 1e6 > a)).sum() for a, b in zip(edges[:-1], edges[1:])]
print("active sequences per 5% time slice:", conc)
top = np.argsort(-dur)[:8]
print("longest:", [(int(L[i]), round(float(dur[i]), 2), round(float((t0[i] - base) / 1e6), 2)) for i in top])
for k in ("steps", "uniq_nodes", "combos", "uniq_w_corr", "uniq_w_scan"):
    print(k, int(c[:, CN.index(k)].sum()))
