"""Build the Turner-2004 parameter tables used by the evaluator (oracle and CUDA).

ViennaRNA's rna_turner2004.par is NOT available on this box, so the
tables are reconstructed:
  * "determined": entries uniquely fixed by the 11,505 (seq, struct, energy) triples the
    reference ships (tools/fit_turner.py -> tools/_fit/determined.json);
  * "symmetry":  int11/int22 partner of a determined entry (loop read from the other strand);
  * "rule":      everything else, filled from the rule the determined entries of the same
    family exhibit (int11: 50+70/AU-GU closure; int21: class values; mismatch*: base+bonus)
    or, for int22, an additive least-squares model rounded to 10 dcal.  These entries are
    PARITY-UNPINNED (no reference artefact touches them); a genuine .par file can be loaded
    at run time through rafft_b200.params.load_par().

Writes rafft_b200/csrc/turner2004.inc (C initialisers) and tools/_fit/provenance.json.
Index conventions: pair types 0..5 = CG GC GU UG AU UA, bases 0..3 = A C G U.
"""
import ast, json, os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
INF = 10000000

HAIRPIN = [INF, INF, INF, 540, 560, 570, 540, 600, 550, 640, 650, 660, 670, 680, 690, 690,
           700, 710, 710, 720, 720, 730, 730, 740, 740, 750, 750, 750, 760, 760, 770]
INTERIOR = [INF, INF, 100, 100, 110, 200, 200, 210, 230, 240, 250, 260, 270, 280, 290, 290,
            300, 310, 310, 320, 330, 330, 340, 340, 350, 350, 350, 360, 360, 370, 370]
BULGE = [INF, 380, 280, 320, 360, 400, 440, 459, 470, 480, 490, 500, 510, 520, 530, 540,
         540, 550, 550, 560, 570, 570, 580, 580, 580, 590, 590, 600, 600, 600, 610]
TRI = {"CAACG": 680, "GUUAC": 690}
TETRA = {"CAACGG": 550, "CCAAGG": 330, "CCACGG": 370, "CCCAGG": 340, "CCGAGG": 350,
         "CCGCGG": 360, "CCUAGG": 370, "CCUCGG": 250, "CUAAGG": 360, "CUACGG": 280,
         "CUCAGG": 370, "CUCCGG": 270, "CUGCGG": 280, "CUUAGG": 350, "CUUCGG": 370,
         "CUUUGG": 370}
HEXA = {"ACAGUACU": 280, "ACAGUGAU": 360, "ACAGUGCU": 290, "ACAGUGUU": 180}
B = "ACGU"


def is_wcgu(a, b):
    return B[a] + B[b] in ("AU", "UA", "CG", "GC", "GU", "UG")


def main():
    d = json.load(open(os.path.join(HERE, "_fit", "determined.json")))
    det = {ast.literal_eval(k): v for k, v in d["determined"].items()}
    prov = {}

    def note(fam, how):
        prov.setdefault(fam, {}).setdefault(how, 0)
        prov[fam][how] += 1

    # ---- scalars
    tAU = det[("tAU",)]
    MLclosing = det[("MLclosing",)]
    assert (tAU, MLclosing) == (50, 930)

    # ---- stack
    stack = np.zeros((6, 6), int)
    for t in range(6):
        for t2 in range(6):
            stack[t, t2] = det[("stack", t, t2)]
            note("stack", "determined")
    assert (stack == stack.T).all()

    # ---- bulge lengths
    bulge = list(BULGE)
    for n in range(1, 31):
        if ("bulge", n) in det:
            bulge[n] = det[("bulge", n)]
            note("bulge", "determined")
        else:
            note("bulge", "rule")

    # ---- dangles
    d5 = np.zeros((6, 4), int)
    d3 = np.zeros((6, 4), int)
    for t in range(6):
        for a in range(4):
            d5[t, a] = det[("d5", t, a)]
            d3[t, a] = det[("d3", t, a)]
            note("dangle", "determined")
            note("dangle", "determined")

    # ---- mismatches
    def mm(name, fill):
        arr = np.zeros((6, 4, 4), int)
        for t in range(6):
            for a in range(4):
                for b in range(4):
                    k = (name, t, a, b)
                    if k in det:
                        arr[t, a, b] = det[k]
                        note(name, "determined")
                    else:
                        arr[t, a, b] = fill(t, a, b)
                        note(name, "rule")
        return arr

    mmH = mm("mmH", lambda t, a, b: 0)
    mmM = mm("mmM", lambda t, a, b: 0)
    mmE = mm("mmE", lambda t, a, b: int(mmM[t, a, b]))

    def base70(t):
        return 0 if t < 2 else 70

    def fill_mmI(t, a, b):
        s = B[a] + B[b]
        return base70(t) + {"AG": -80, "GA": -100, "GG": -100, "UU": -60}.get(s, 0)

    def fill_mm23(t, a, b):
        s = B[a] + B[b]
        bonus = {"GA": -110 if t != 1 else -120, "GG": -70, "UU": -30}.get(s, 0)
        if s == "AG":
            bonus = -50 if t in (0, 3, 5) else 0
        return base70(t) + bonus

    mmI = mm("mmI", fill_mmI)
    mm1n = mm("mm1n", lambda t, a, b: base70(t))
    mm23 = mm("mm23", fill_mm23)
    # sanity: the fill rules must reproduce every determined non-WC entry of their family
    for name, arr, f in (("mmI", mmI, fill_mmI), ("mm1n", mm1n, lambda t, a, b: base70(t)),
                         ("mm23", mm23, fill_mm23)):
        bad = [(t, a, b, int(arr[t, a, b]), f(t, a, b)) for t in range(6) for a in range(4)
               for b in range(4) if (name, t, a, b) in det and arr[t, a, b] != f(t, a, b)]
        print(name, "rule mismatches on determined entries:", bad)

    # ---- int11
    int11 = np.full((6, 6, 4, 4), INF, int)
    for k, v in det.items():
        if k[0] == "int11":
            int11[k[1:]] = v
    src = int11.copy()
    for t in range(6):
        for t2 in range(6):
            for a in range(4):
                for b in range(4):
                    if src[t, t2, a, b] != INF:
                        note("int11", "determined")
                    elif src[t2, t, b, a] != INF:
                        int11[t, t2, a, b] = src[t2, t, b, a]
                        note("int11", "symmetry")
                    else:
                        v = 50 + base70(t) + base70(t2)
                        if B[a] + B[b] == "GG":
                            v = -140 if (t < 2 or t2 < 2) else -70
                        int11[t, t2, a, b] = v
                        note("int11", "rule")
    bad = sum(1 for t in range(6) for t2 in range(6) for a in range(4) for b in range(4)
              if src[t, t2, a, b] != INF and src[t2, t, b, a] != INF
              and src[t, t2, a, b] != src[t2, t, b, a])
    print("int11 symmetry violations among determined:", bad)

    # ---- int21: class rule on (number of AU/GU closures, mismatch class)
    int21 = np.full((6, 6, 4, 4, 4), INF, int)
    for k, v in det.items():
        if k[0] == "int21":
            int21[k[1:]] = v
    GEN = [230, 300, 370]
    PUR = [110, 190, 260]
    UU = [150, 220, 300]

    def rule21(t, t2, a, b, c):
        k = (t >= 2) + (t2 >= 2)
        pairs = (B[a] + B[b], B[a] + B[c])
        if any(p in ("AG", "GA", "GG") for p in pairs):
            return PUR[k]
        if any(p == "UU" for p in pairs):
            return UU[k]
        return GEN[k]
    nbad = 0
    for t in range(6):
        for t2 in range(6):
            for a in range(4):
                for b in range(4):
                    for c in range(4):
                        if int21[t, t2, a, b, c] != INF:
                            note("int21", "determined")
                            if int21[t, t2, a, b, c] != rule21(t, t2, a, b, c):
                                nbad += 1
                        else:
                            int21[t, t2, a, b, c] = rule21(t, t2, a, b, c)
                            note("int21", "rule")
    print("int21 determined entries that deviate from the class rule (kept as determined):", nbad)

    # ---- int22: determined + symmetry + additive LSQ model
    int22 = np.full((6, 6, 4, 4, 4, 4), INF, int)
    for k, v in det.items():
        if k[0] == "int22":
            int22[k[1:]] = v
    src = int22.copy()
    # model: v = h[t][a][d] + h[t2][c][b] + D[cls(a,d)][cls(c,b)]  (second end read from the
    # other strand: closing pair (q,p), 5' neighbour sq1=c, 3' neighbour sp1=b)
    CL = {"AG": 1, "GA": 2, "GG": 3, "UU": 4}

    def cls(a, b):
        s = B[a] + B[b]
        if is_wcgu(a, b):
            return 5
        return CL.get(s, 0)
    rows, rhs = [], []
    nh = 6 * 16
    ncl = 6

    def feat(t, t2, a, b, c, dd):
        f = np.zeros(nh + ncl * ncl)
        f[t * 16 + a * 4 + dd] += 1
        f[t2 * 16 + c * 4 + b] += 1
        c1, c2 = cls(a, dd), cls(c, b)
        lo, hi = min(c1, c2), max(c1, c2)
        f[nh + lo * ncl + hi] += 1
        return f
    for k, v in det.items():
        if k[0] == "int22":
            rows.append(feat(*k[1:]))
            rhs.append(v)
    A = np.array(rows)
    x = np.linalg.lstsq(A, np.array(rhs, float), rcond=None)[0]
    pred = A @ x
    exact = int((np.round(pred / 10) * 10 == np.array(rhs)).sum())
    print("int22 additive model: %d/%d determined entries reproduced exactly, rms %.1f dcal"
          % (exact, len(rhs), float(np.sqrt(((pred - rhs) ** 2).mean()))))
    bad = 0
    for t in range(6):
        for t2 in range(6):
            for a in range(4):
                for b in range(4):
                    for c in range(4):
                        for dd in range(4):
                            if src[t, t2, a, b, c, dd] != INF:
                                note("int22", "determined")
                                s = src[t2, t, c, dd, a, b]
                                if s != INF and s != src[t, t2, a, b, c, dd]:
                                    bad += 1
                            elif src[t2, t, c, dd, a, b] != INF:
                                int22[t, t2, a, b, c, dd] = src[t2, t, c, dd, a, b]
                                note("int22", "symmetry")
                            else:
                                v = float(feat(t, t2, a, b, c, dd) @ x)
                                int22[t, t2, a, b, c, dd] = int(round(v / 10.0)) * 10
                                note("int22", "rule")
    print("int22 symmetry violations among determined:", bad)
    # make the rule-filled part symmetric as well
    for t in range(6):
        for t2 in range(6):
            for a in range(4):
                for b in range(4):
                    for c in range(4):
                        for dd in range(4):
                            if src[t, t2, a, b, c, dd] == INF and src[t2, t, c, dd, a, b] == INF:
                                if (t, t2, a, b, c, dd) > (t2, t, c, dd, a, b):
                                    int22[t, t2, a, b, c, dd] = int22[t2, t, c, dd, a, b]

    # ---- touched-but-undetermined keys: the rule fill above must not break the recorded
    # energies.  Solve A_U * delta = residual for the smallest (L1) integer correction.
    z = np.load(os.path.join(HERE, "_fit", "system.npz"))
    A = z["A"].astype(float)
    bvec = z["b"].astype(float)
    keys = [ast.literal_eval(k) for k in z["keys"]]
    und = set(ast.literal_eval(k) for k in d["undetermined"])
    fam = {"stack": stack, "d5": d5, "d3": d3, "mmH": mmH, "mmM": mmM, "mmE": mmE, "mmI": mmI,
           "mm1n": mm1n, "mm23": mm23, "int11": int11, "int21": int21, "int22": int22}
    sp_all = dict(TRI); sp_all.update(TETRA); sp_all.update(HEXA)

    def lookup(k):
        if k[0] == "tAU": return tAU
        if k[0] == "MLclosing": return MLclosing
        if k[0] == "bulge": return bulge[k[1]]
        if k[0] == "sp": return sp_all[k[1]]
        return int(fam[k[0]][k[1:]])
    x0 = np.array([lookup(k) for k in keys], float)
    r = bvec - A @ x0
    print("rows with non-zero residual before correction:", int((np.abs(r) > 1e-6).sum()))
    uidx = [i for i, k in enumerate(keys) if k in und]
    rows = np.where(np.abs(A[:, uidx]).sum(1) > 0)[0]
    assert set(np.where(np.abs(r) > 1e-6)[0]) <= set(rows)
    from scipy.optimize import linprog
    AU = A[np.ix_(rows, uidx)]
    nU = len(uidx)
    res = linprog(np.ones(2 * nU), A_eq=np.hstack([AU, -AU]), b_eq=r[rows],
                  bounds=[(0, None)] * (2 * nU), method="highs-ds")
    assert res.status == 0, res.message
    delta = res.x[:nU] - res.x[nU:]
    assert np.abs(delta - np.round(delta)).max() < 1e-6, "non-integer correction"
    nchg = 0
    for i, dv in zip(uidx, np.round(delta).astype(int)):
        if dv:
            k = keys[i]
            fam[k[0]][k[1:]] += dv
            nchg += 1
    x1 = np.array([lookup(k) for k in keys], float)
    print("undetermined keys adjusted:", nchg, " max |residual| after:", np.abs(bvec - A @ x1).max())
    odd = [(keys[i], lookup(keys[i])) for i in uidx if lookup(keys[i]) % 10]
    print("adjusted values that are not multiples of 10:", odd)
    for k in und:
        prov[k[0]]["rule"] -= 1
        prov[k[0]].setdefault("constrained", 0)
        prov[k[0]]["constrained"] += 1

    # ---- specials
    for name, tab in (("tri", TRI), ("tetra", TETRA), ("hexa", HEXA)):
        for s in tab:
            if ("sp", s) in det:
                assert det[("sp", s)] == tab[s], (s, det[("sp", s)], tab[s])
                note("special_hairpins", "determined")
            else:
                note("special_hairpins", "rule")

    # ---- emit
    def carr(a):
        a = np.asarray(a)
        return ",".join(str(int(v)) for v in a.ravel())

    def cstr(tab, n):
        return "{" + ",".join('{"%s",%d}' % (s, tab[s]) for s in sorted(tab)) + "}"
    out = os.path.join(ROOT, "rafft_b200", "csrc", "turner2004.inc")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    with open(out, "w") as f:
        f.write("/* GENERATED by tools/gen_tables.py -- Turner-2004 nearest-neighbour parameters in\n"
                " * dcal/mol, reconstructed from the reference's recorded ViennaRNA energies\n"
                " * (benchmark_results/*_scores.csv; provenance in tools/_fit/provenance.json).\n"
                " * Pair types 0..5 = CG GC GU UG AU UA, bases 0..3 = A C G U. */\n")
        f.write("#define T04_INF %d\n" % INF)
        f.write("#define T04_TERMINAL_AU %d\n#define T04_ML_CLOSING %d\n" % (tAU, MLclosing))
        f.write("#define T04_ML_INTERN -90\n#define T04_ML_BASE 0\n")
        f.write("#define T04_NINIO 60\n#define T04_MAX_NINIO 300\n#define T04_LXC 107.856\n")
        f.write("#define T04_STACK {%s}\n" % carr(stack))
        f.write("#define T04_HAIRPIN {%s}\n" % carr(HAIRPIN))
        f.write("#define T04_BULGE {%s}\n" % carr(bulge))
        f.write("#define T04_INTERIOR {%s}\n" % carr(INTERIOR))
        f.write("#define T04_MISMATCH_H {%s}\n" % carr(mmH))
        f.write("#define T04_MISMATCH_I {%s}\n" % carr(mmI))
        f.write("#define T04_MISMATCH_1NI {%s}\n" % carr(mm1n))
        f.write("#define T04_MISMATCH_23I {%s}\n" % carr(mm23))
        f.write("#define T04_MISMATCH_M {%s}\n" % carr(mmM))
        f.write("#define T04_MISMATCH_EXT {%s}\n" % carr(mmE))
        f.write("#define T04_DANGLE5 {%s}\n" % carr(d5))
        f.write("#define T04_DANGLE3 {%s}\n" % carr(d3))
        f.write("#define T04_INT11 {%s}\n" % carr(int11))
        f.write("#define T04_INT21 {%s}\n" % carr(int21))
        f.write("#define T04_INT22 {%s}\n" % carr(int22))
        f.write("#define T04_N_TRI %d\n#define T04_TRILOOPS %s\n" % (len(TRI), cstr(TRI, 5)))
        f.write("#define T04_N_TETRA %d\n#define T04_TETRALOOPS %s\n" % (len(TETRA), cstr(TETRA, 6)))
        f.write("#define T04_N_HEXA %d\n#define T04_HEXALOOPS %s\n" % (len(HEXA), cstr(HEXA, 8)))
    json.dump(prov, open(os.path.join(HERE, "_fit", "provenance.json"), "w"), indent=1, sort_keys=True)
    print(json.dumps(prov, sort_keys=True))
    print("wrote", out)


if __name__ == "__main__":
    main()
