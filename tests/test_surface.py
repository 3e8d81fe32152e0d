"""Landscape distance matrix (SURVEY.md 8(f) "next" #4; reference utility/surface.py:19-40).

not gpu: the oracle restatement of RNA.bp_distance (oracle/surface_oracle.py; parity unpinned, see
         its header) on hand-counted cases and its metric properties; the parser.
gpu:     rafft_bp_distance_matrix through the C-ABI, bit-exact against the oracle."""
import os
import random
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
GOLD = os.path.join(ROOT, "tests", "golden")


def _random_structure(rng, L, p_open=0.3):
    "a random balanced round-bracket structure with hairpins of >= 3 unpaired positions"
    out = ["."] * L

    def fill(i, j):
        while j - i >= 4:
            if rng.random() < p_open:
                a = rng.randrange(i, j - 3)
                b = rng.randrange(a + 4, j + 1)
                out[a], out[b] = "(", ")"
                fill(a + 1, b - 1)
                i = b + 1
            else:
                i += 1
    fill(0, L - 1)
    return "".join(out)


def test_bp_distance_hand_counted():
    import surface_oracle as so
    assert so.bp_distance("((...))", "((...))") == 0
    assert so.bp_distance("((...))", ".(...).") == 1
    assert so.bp_distance("((...))", ".......") == 2
    assert so.bp_distance("(....).", ".(....)") == 2
    assert so.bp_distance("((...))..(...)", "(.(...).....).") == 5      # no common pair: 3 + 2
    assert so.pair_table("<[.>].()") == [-1, -1, -1, -1, -1, -1, 7, 6]  # only round brackets pair


def test_bp_distance_is_symmetric_difference_of_pair_sets():
    import surface_oracle as so
    from kin_oracle import paired_positions
    rng = random.Random(5)
    for _ in range(200):
        L = rng.randrange(5, 90)
        a, b, c = (_random_structure(rng, L) for _ in range(3))
        pa, pb = set(paired_positions(a)), set(paired_positions(b))
        assert so.bp_distance(a, b) == len(pa ^ pb) == so.bp_distance(b, a)
        assert so.bp_distance(a, c) <= so.bp_distance(a, b) + so.bp_distance(b, c)


def test_surface_parser_keeps_first_occurrences():
    import surface_oracle as so
    from rafft_b200 import surface
    path = os.path.join(GOLD, "rafft_20.out")
    res, seq = surface.parse_rafft_output(path)
    assert (res, seq) == so.parse_rafft_output(path)
    assert len({s for s, _ in res}) == len(res) > 20 and len(seq) == len(res[0][0])
    assert res[0][0] == "." * len(seq)


@pytest.mark.gpu
def test_distance_matrix_vs_oracle_example():
    import surface_oracle as so
    from rafft_b200 import surface
    res, _ = surface.parse_rafft_output(os.path.join(GOLD, "rafft_20.out"))
    got = surface.get_distance_matrix(res)
    want = np.array(so.get_distance_matrix(res))
    assert got.dtype == np.float64 and got.shape == want.shape
    assert (got == want).all()
    assert (got == got.T).all() and (np.diag(got) == 0).all()


@pytest.mark.gpu
@pytest.mark.parametrize("n,L", [(1, 7), (2, 1), (31, 33), (33, 127), (64, 128), (97, 129), (130, 300), (40, 3000)])
def test_distance_matrix_vs_oracle_random(n, L):
    import surface_oracle as so
    from rafft_b200 import surface
    rng = random.Random(1000 * n + L)
    structs = [(_random_structure(rng, L), 0.0) for _ in range(n)]
    got = surface.get_distance_matrix(structs)
    want = np.array(so.get_distance_matrix(structs)).reshape(n, n)
    assert (got == want).all()


@pytest.mark.gpu
def test_distance_matrix_other_brackets_and_errors():
    import surface_oracle as so
    from rafft_b200 import surface
    structs = [("<[.>].()", 0.0), ("(......)", 0.0), ("........", 0.0)]
    got = surface.get_distance_matrix(structs)
    assert (got == np.array(so.get_distance_matrix(structs))).all()
    with pytest.raises(ValueError):
        surface.get_distance_matrix([("..)", 0.0), ("...", 0.0)])
    with pytest.raises(ValueError):
        surface.get_distance_matrix([("....", 0.0), ("...", 0.0)])
    assert surface.get_distance_matrix([]).shape == (0, 0)


@pytest.mark.gpu
def test_fold_then_distance_matrix_cfg5():
    "trajectory of a 200-nt fold -> unique structures -> distance matrix, against the oracle"
    import surface_oracle as so
    from rafft_b200 import fold, surface
    from rafft_b200.synth import splitmix_seq
    seq = splitmix_seq(200, 200)
    _, traj = fold(seq, nb_mode=100, max_stack=60, max_branch=1000, traj=True)
    seen, res = set(), []
    for step in traj:
        for st in step:
            if st.str_struct not in seen:
                seen.add(st.str_struct)
                res.append((st.str_struct, st.energy))
    got = surface.get_distance_matrix(res)
    want = np.array(so.get_distance_matrix(res))
    assert (got == want).all() and len(res) > 100
