"""Per-iteration, per-kernel-family device times of one fold (profile mode: CUDA events around
every launch):  python tools/iter_profile.py cfg2|cfg3|cfg4 [count]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
os.environ["RAFFT_DEBUG_ALLOC"] = "1"
import numpy as np
from rafft_b200 import rafft as R
from rafft_b200.synth import config_batch
wl = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
count = int(sys.argv[2]) if len(sys.argv) > 2 else None
n, ms = (100, 50) if wl == "cfg2" else (200, 200)
seqs = config_batch("cfg4" if wl == "cfg4" else "cfg2", count)
ctx = R.get_context(0)
offs = np.zeros(len(seqs) + 1, dtype=np.int64); np.cumsum([len(s) for s in seqs], out=offs[1:])
prm = R.make_params(n, ms, 1000, traj=False)
sb = "".join(seqs).encode()
for rep in range(2):
    ctx.set_option("profile", 0)
    raw = ctx.fold_raw(sb, offs, prm); raw.close()
ctx.set_option("profile", 1)
raw = ctx.fold_raw(sb, offs, prm)
print("families (ms):", {k: round(v, 3) for k, v in raw.family_ms.items()}, "kernel_ms", round(raw.kernel_ms, 3), file=sys.stderr)
