"""Wall time of consecutive rafft_fold_batch calls (host buffers) on a benchmark workload."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
from rafft_b200 import rafft as R
from rafft_b200.synth import config_batch
wl = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
n, ms = (100, 50) if wl == "cfg2" else (200, 200)
seqs = config_batch("cfg2")
ctx = R.get_context(0)
offs = np.zeros(len(seqs) + 1, dtype=np.int64); np.cumsum([len(s) for s in seqs], out=offs[1:])
prm = R.make_params(n, ms, 1000, traj=False)
sb = "".join(seqs).encode()
keep = None
for it in range(6):
    t = time.perf_counter()
    raw = ctx.fold_raw(sb, offs, prm)
    dt = time.perf_counter() - t
    print("call %d: wall %.1f ms kernel %.1f ms d2h %.0f MB" % (it, dt * 1e3, raw.kernel_ms, raw.d2h_bytes / 1e6), flush=True)
    if it % 2 == 0: keep = raw
    else: raw.close()
