import gzip, json, sys, numpy as np
sys.path.insert(0, "/root/repo")
from rafft_b200 import rafft_kin as K
class S:
    def __init__(s, db, e): s.str_struct, s.energy = db, e
g = json.load(gzip.open("/root/repo/tests/golden/kin_golden.json.gz", "rt"))
for case in g:
    paths = [[S(db, e) for db, e in st] for st in case["steps"]]
    traj, times, sl, equi = K.kinetics(paths, case["max_time"], case["n_steps"])
    for r, want in zip(case["traj_rows"], case["traj"]):
        got = np.asarray(traj[r], dtype=np.float64); want = np.asarray(want)
        print(case["name"], r, "max abs %.3e" % np.abs(got - want).max(), "L1 %.3e" % np.abs(got - want).sum(), "sum got %.6f min got %.3e min want %.3e" % (got.sum(), got.min(), want.min()))
