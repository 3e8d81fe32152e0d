"""ORACLE -- test infrastructure, NOT the product.  Only tests/ may import this.

Restatement of the landscape distance matrix of the reference's utility/surface.py:19-26
(get_distance_matrix: RNA.bp_distance for every pair i < j, mirrored, zero diagonal, float matrix)
and of its parser (parse_rafft_output, :29-40).

RNA.bp_distance is third-party arithmetic that is absent from /root/reference: ViennaRNA (README.org:8
asks for >= 2.4.17; rafft-rs/.github/build_env.yml:8 pins 2.4.18), `vrna_bp_distance` in
src/ViennaRNA/utils/structure_utils.c.  Its published algorithm: both dot-bracket strings become
1-based pair tables (only round brackets pair); over i = 1..min(len): where the tables differ, add
one if structure 1 pairs i with a later position and one if structure 2 does -- the number of base
pairs present in exactly one of the two structures.

PARITY UNPINNED for this row: the reference holds no distance-matrix fixture and ViennaRNA cannot
be imported here; the restatement is anchored on the definition above and on its metric properties
(tests/test_surface.py)."""


def pair_table(structure):
    "0-based partner or -1; only ( and ) pair (vrna_ptable)"
    pt = [-1] * len(structure)
    pile = []
    for i, c in enumerate(structure):
        if c == "(":
            pile.append(i)
        elif c == ")":
            j = pile.pop()
            pt[i], pt[j] = j, i
    return pt


def bp_distance(s1, s2):
    t1, t2 = pair_table(s1), pair_table(s2)
    dist = 0
    for i in range(min(len(s1), len(s2))):
        if t1[i] != t2[i]:
            if t1[i] > i:
                dist += 1
            if t2[i] > i:
                dist += 1
    return dist


def get_distance_matrix(structures):
    "structures: [(dot-bracket, energy)] -> list of lists of floats (utility/surface.py:19-26)"
    n = len(structures)
    matrix = [[0.0] * n for _ in range(n)]
    for si in range(n):
        for sj in range(si + 1, n):
            d = float(bp_distance(structures[si][0], structures[sj][0]))
            matrix[si][sj] = d
            matrix[sj][si] = d
    return matrix


def parse_rafft_output(infile):
    results, seen = [], set()
    with open(infile) as fh:
        seq = fh.readline().strip()
        for line in fh:
            if not line.startswith("#"):
                struct, nrj = line.strip().split()
                if struct not in seen:
                    results.append((struct, float(nrj)))
                    seen.add(struct)
    return results, seq
