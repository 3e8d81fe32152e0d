"""Recover Turner-2004 table entries from the reference's recorded ViennaRNA energies.

Reads the (seq, struct, nrj) triples in /root/reference/benchmark_results/*_scores.csv
(nrj = RNA.energy_of_struct, benchmark_results/scoring.py:125), decomposes each with
tools/loopdecomp.py and solves the linear system.  Writes tools/_fit/determined.json with
every key whose value is uniquely determined (exact multiple of 1 dcal) and the set of
touched-but-undetermined keys.  Run here only (needs /root/reference); the output feeds
tools/gen_tables.py.
"""
import csv, glob, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(__file__))
from loopdecomp import decompose

REF = "/root/reference/benchmark_results"
OUT = os.path.join(os.path.dirname(__file__), "_fit")


def load_triples():
    tri = {}
    for f in sorted(glob.glob(REF + "/*_scores.csv")):
        for r in csv.DictReader(open(f)):
            seq, st = r["seq"], r["struct"]
            if len(seq) != len(st) or set(seq) - set("ACGU") or set(st) - set("()."):
                continue
            tri[(seq, st)] = int(round(float(r["nrj"]) * 100))
    return tri


def main():
    os.makedirs(OUT, exist_ok=True)
    tri = load_triples()
    print("triples", len(tri))
    rows, rhs, kidx = [], [], {}
    for (seq, st), e in tri.items():
        try:
            keys, const = decompose(seq, st)
        except KeyError:
            continue
        rows.append({kidx.setdefault(k, len(kidx)): n for k, n in keys.items()})
        rhs.append(e - const)
    A = np.zeros((len(rows), len(kidx)))
    for r, d in enumerate(rows):
        for c, n in d.items():
            A[r, c] = n
    b = np.array(rhs, float)
    print("system", A.shape)
    # normal equations eigen-analysis (fast: 3k x 3k)
    AtA = A.T @ A
    w, V = np.linalg.eigh(AtA)
    tol = 1e-6  # gap: null eigenvalues ~1e-15, smallest real one ~1e-2
    nullmask = w < tol
    print("rank", int((~nullmask).sum()), "nullity", int(nullmask.sum()))
    Vn = V[:, nullmask]
    Vr = V[:, ~nullmask]
    x = Vr @ ((Vr.T @ (A.T @ b)) / w[~nullmask])
    res = A @ x - b
    print("max |residual|", np.abs(res).max())
    det = (np.abs(Vn) ** 2).sum(1) < 1e-9
    keys = sorted(kidx, key=lambda k: kidx[k])
    determined, undetermined = {}, []
    bad = 0
    for k, i in kidx.items():
        if det[i]:
            v = x[i]
            if abs(v - round(v)) > 1e-3:
                bad += 1
            determined[repr(k)] = int(round(v))
        else:
            undetermined.append(repr(k))
    print("determined", len(determined), "undetermined", len(undetermined), "non-integer", bad)
    json.dump({"determined": determined, "undetermined": undetermined},
              open(OUT + "/determined.json", "w"), indent=0)
    # keep the system for gauge-resolution by gen_tables
    np.savez_compressed(OUT + "/system.npz", A=A.astype(np.int16), b=b.astype(np.int32),
                        keys=np.array([repr(k) for k in keys]))


if __name__ == "__main__":
    main()
