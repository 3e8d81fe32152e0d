import sys, time
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/oracle')
import pyoracle
import rafft_b200
from rafft_b200 import rafft as R
from rafft_b200.synth import splitmix_seq, config_batch
ctx = R.get_context(0)
import numpy as np
def run(seqs, **kw):
    offs = np.zeros(len(seqs)+1, dtype=np.int64); np.cumsum([len(s) for s in seqs], out=offs[1:])
    return ctx.fold_raw("".join(seqs).encode(), offs, R.make_params(traj=True, **kw))
seqs = [splitmix_seq(1,120), splitmix_seq(7, 300), "GGGAAACCC", splitmix_seq(3, 60)]
t=time.time(); raw = run(seqs, nb_mode=100, max_stack=50, max_branch=1000); print("gpu", time.time()-t, raw.kernel_ms, raw.status)
for i,s in enumerate(seqs):
    o = pyoracle.fold(s, 100, 50, 1000)
    g = raw.steps(i)
    print(i, len(s), "steps", len(o), len(g), "MATCH" if o==g else "DIFF")
    print(dict(zip(R.N.COUNTER_NAMES, raw.counters[i])))
b = config_batch('cfg2', 512)
t=time.time(); raw = run(b, nb_mode=100, max_stack=50, max_branch=1000); print("gpu512", time.time()-t, raw.kernel_ms, np.bincount(raw.status))
t=time.time(); o,_ = pyoracle.fold_batch(b, 100, 50, 1000); print("oracle512", time.time()-t)
bad=[i for i in range(len(b)) if raw.steps(i)!=o[i]]
print("mismatches", len(bad), bad[:10])
