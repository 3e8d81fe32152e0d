"""Generate tests/golden/kin_golden.json.gz with the UNMODIFIED reference (build container only):
for a few trajectory files, what rafft.rafft_kin computes -- connectivity per step
(get_connected_prev, rafft_kin.py:48-56), the unique structure list, and the populations of
kinetics() (rafft_kin.py:94-150).  The cfg5 trajectory (200 nt, -ms 200) is folded by the oracle
(pinned to the reference, tests/test_oracle.py) because the reference's Python fold takes long."""
import gzip, json, os, sys
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
import refrun  # noqa: F401  (puts /root/reference and the shims on sys.path)
from refrun import splitmix_seq
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import pyoracle
from rafft.rafft_kin import kinetics, get_connected_prev
from rafft.utils import parse_rafft_output
import tempfile


def case(name, path, max_time, n_steps):
    fast_paths, seq = parse_rafft_output(path)
    conn = [[get_connected_prev(st, fast_paths[t - 1]) for st in step] for t, step in enumerate(fast_paths)]
    traj, times, struct_list, equi = kinetics(fast_paths, max_time, n_steps)
    rows = [0, 1, n_steps // 4, n_steps // 2, n_steps]
    return {"name": name, "seq": seq, "max_time": max_time, "n_steps": n_steps,
            "steps": [[(st.str_struct, st.energy) for st in step] for step in fast_paths],
            "conn": conn, "struct_list": [st.str_struct for st in struct_list],
            "equi": [(s, float(e), float(p), int(i)) for s, e, p, i in equi],
            "times": [float(t) for t in times], "traj_rows": rows,
            "traj": [[float(x) for x in traj[r]] for r in rows]}


def main():
    out = []
    g = os.path.join(ROOT, "tests", "golden")
    out.append(case("rafft.out", os.path.join(g, "rafft.out"), 30, 100))
    out.append(case("rafft_20.out", os.path.join(g, "rafft_20.out"), 40, 100))
    # cfg5: one 200-nt sequence (seed 200), -n 100 -ms 200 --traj
    seq = splitmix_seq(200, 200)
    steps = pyoracle.fold(seq, 100, 200, 1000)
    with tempfile.NamedTemporaryFile("w", suffix=".out", delete=False) as fh:
        fh.write(seq + "\n")
        for si, step in enumerate(steps):
            fh.write("# {:-^20}\n".format(si))
            for db, e in step:
                import numpy as np
                fh.write("%s %6.1f\n" % (db, float(np.float32(e / 100.0))))
        path = fh.name
    out.append(case("cfg5", path, 30, 100))
    os.unlink(path)
    with gzip.open(os.path.join(g, "kin_golden.json.gz"), "wt") as fh:
        json.dump(out, fh)
    for c in out:
        print(c["name"], "steps", len(c["steps"]), "structures", sum(map(len, c["steps"])), "unique", len(c["struct_list"]))


if __name__ == "__main__":
    main()
