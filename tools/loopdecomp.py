"""Loop decomposition of an RNA secondary structure into Turner-2004 parameter keys.

Tooling for tools/fit_turner.py (table reconstruction) — NOT on the product path.
Follows the ViennaRNA default model (dangles=2, special hairpins on) as restated in
SURVEY.md Appendix B; call sites in the reference: rafft/utils.py:135-138,
benchmark_results/scoring.py:125.

Every structure maps to {key: count} + a constant (terms taken from fixed length
arrays).  Pair types 0..5 = CG GC GU UG AU UA; bases 0..3 = A C G U.
"""
import math

B = {"A": 0, "C": 1, "G": 2, "U": 3}
PT = {(1, 2): 0, (2, 1): 1, (2, 3): 2, (3, 2): 3, (0, 3): 4, (3, 0): 5}

HAIRPIN = [None, None, None, 540, 560, 570, 540, 600, 550, 640, 650, 660, 670, 680, 690, 690,
           700, 710, 710, 720, 720, 730, 730, 740, 740, 750, 750, 750, 760, 760, 770]
INTERIOR = [None, None, 100, 100, 110, 200, 200, 210, 230, 240, 250, 260, 270, 280, 290, 290,
            300, 310, 310, 320, 330, 330, 340, 340, 350, 350, 350, 360, 360, 370, 370]
BULGE = [None, 380, 280, 320, 360, 400, 440, 459, 470, 480, 490, 500, 510, 520, 530, 540,
         540, 550, 550, 560, 570, 570, 580, 580, 580, 590, 590, 600, 600, 600, 610]
LXC = 107.856
NINIO, MAX_NINIO = 60, 300
ML_INTERN, ML_BASE = -90, 0
TRILOOPS = {"CAACG", "GUUAC"}
TETRALOOPS = {"CAACGG", "CCAAGG", "CCACGG", "CCCAGG", "CCGAGG", "CCGCGG", "CCUAGG", "CCUCGG",
              "CUAAGG", "CUACGG", "CUCAGG", "CUCCGG", "CUGCGG", "CUUAGG", "CUUCGG", "CUUUGG"}
HEXALOOPS = {"ACAGUACU", "ACAGUGAU", "ACAGUGCU", "ACAGUGUU"}


def pair_table(struct):
    pt = [-1] * len(struct)
    st = []
    for i, c in enumerate(struct):
        if c == "(":
            st.append(i)
        elif c == ")":
            j = st.pop()
            pt[j], pt[i] = i, j
    return pt


def loglen(arr, u):
    return arr[u] if u <= 30 else arr[30] + int(LXC * math.log(u / 30.0))


def decompose(seq, struct, free_bulge=True):
    """-> (dict key->count, const)"""
    s = [B[c] for c in seq]
    L = len(s)
    pt = pair_table(struct)
    keys, const = {}, 0

    def add(k, n=1):
        keys[k] = keys.get(k, 0) + n

    def ty(i, j):
        return PT[(s[i], s[j])]

    # exterior loop
    p = 0
    while p < L:
        if pt[p] > p:
            i, j = p, pt[p]
            t = ty(i, j)
            if i > 0 and j < L - 1:
                add(("mmE", t, s[i - 1], s[j + 1]))
            elif i > 0:
                add(("d5", t, s[i - 1]))
            elif j < L - 1:
                add(("d3", t, s[j + 1]))
            if t >= 2:
                add(("tAU",))
            p = j + 1
        else:
            p += 1
    # closed loops
    for i in range(L):
        j = pt[i]
        if j <= i:
            continue
        t = ty(i, j)
        br = []
        p = i + 1
        while p < j:
            if pt[p] > p:
                br.append((p, pt[p]))
                p = pt[p] + 1
            else:
                p += 1
        if not br:
            u = j - i - 1
            loop = seq[i:j + 1]
            if (u == 3 and loop in TRILOOPS) or (u == 4 and loop in TETRALOOPS) or \
               (u == 6 and loop in HEXALOOPS):
                add(("sp", loop))
                continue
            const += loglen(HAIRPIN, u)
            if u == 3:
                if t >= 2:
                    add(("tAU",))
            else:
                add(("mmH", t, s[i + 1], s[j - 1]))
        elif len(br) == 1:
            p, q = br[0]
            t2 = ty(q, p)
            n1, n2 = p - i - 1, j - q - 1
            nl, ns = max(n1, n2), min(n1, n2)
            if nl == 0:
                add(("stack", t, t2))
            elif ns == 0:
                if free_bulge and nl <= 30:
                    add(("bulge", nl))
                else:
                    const += loglen(BULGE, nl)
                if nl == 1:
                    add(("stack", t, t2))
                else:
                    if t >= 2:
                        add(("tAU",))
                    if t2 >= 2:
                        add(("tAU",))
            elif ns == 1:
                if nl == 1:
                    add(("int11", t, t2, s[i + 1], s[j - 1]))
                elif nl == 2:
                    if n1 == 1:
                        add(("int21", t, t2, s[i + 1], s[q + 1], s[j - 1]))
                    else:
                        add(("int21", t2, t, s[q + 1], s[i + 1], s[p - 1]))
                else:
                    const += loglen(INTERIOR, nl + 1) + min(MAX_NINIO, (nl - 1) * NINIO)
                    add(("mm1n", t, s[i + 1], s[j - 1]))
                    add(("mm1n", t2, s[q + 1], s[p - 1]))
            elif ns == 2 and nl == 2:
                add(("int22", t, t2, s[i + 1], s[p - 1], s[q + 1], s[j - 1]))
            elif ns == 2 and nl == 3:
                const += INTERIOR[5] + NINIO
                add(("mm23", t, s[i + 1], s[j - 1]))
                add(("mm23", t2, s[q + 1], s[p - 1]))
            else:
                const += loglen(INTERIOR, n1 + n2) + min(MAX_NINIO, (nl - ns) * NINIO)
                add(("mmI", t, s[i + 1], s[j - 1]))
                add(("mmI", t2, s[q + 1], s[p - 1]))
        else:
            add(("MLclosing",))
            unp = j - i - 1
            tt = ty(j, i)
            const += ML_INTERN
            add(("mmM", tt, s[j - 1], s[i + 1]))
            if tt >= 2:
                add(("tAU",))
            for p, q in br:
                unp -= q - p + 1
                tt = ty(p, q)
                const += ML_INTERN
                add(("mmM", tt, s[p - 1], s[q + 1]))
                if tt >= 2:
                    add(("tAU",))
            const += ML_BASE * unp
    return keys, const
