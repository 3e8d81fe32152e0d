"""Kinetics on the folding graph (SURVEY.md 8(f) "next" #1; reference rafft/rafft_kin.py).

not gpu: the oracle restatement (oracle/kin_oracle.py) against golden vectors produced by the
         unmodified reference (tools/make_golden_kin.py); host logic of the mirror.
gpu:     rafft_kin_graph through the C-ABI against the oracle (bit-exact connectivity and
         numbering) and kinetics() populations against the reference's within 1e-9."""
import gzip
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
GOLD = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="module")
def kin_golden():
    with gzip.open(os.path.join(GOLD, "kin_golden.json.gz"), "rt") as fh:
        return json.load(fh)


class _S:
    def __init__(self, db, e):
        self.str_struct, self.energy = db, e


def _paths(case):
    return [[_S(db, e) for db, e in step] for step in case["steps"]]


def test_kin_oracle_matches_reference(kin_golden):
    import kin_oracle
    for case in kin_golden:
        steps = [[db for db, _ in step] for step in case["steps"]]
        uniq, conn = kin_oracle.folding_graph(steps)
        assert conn == case["conn"], case["name"]
        flat = [s for st in steps for s in st]
        assert [flat[i] for i in range(len(flat)) if uniq[i] == i] == case["struct_list"]


def test_kin_oracle_brackets():
    import kin_oracle
    assert kin_oracle.paired_positions("((..))") == [(1, 4), (0, 5)]
    assert kin_oracle.paired_positions("<[.>].(") == [(0, 3), (1, 4)]   # pk pile, unmatched opener ignored
    with pytest.raises(IndexError):
        kin_oracle.paired_positions("..)")


def test_kin_cli_flags():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bin", "rafft_kin"), "--help"], capture_output=True, text=True)
    assert out.returncode == 0
    for flag in ("--n_steps", "--max_time", "--init_pop", "--show_thres", "--plot", "--uni", "--other_rate"):
        assert flag in out.stdout


# ------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_kin_graph_vs_oracle(ctx, kin_golden):
    import kin_oracle
    from rafft_b200 import rafft_kin as K
    for case in kin_golden:
        paths = _paths(case)
        uniq, conn, conn_off, step_off = K.folding_graph(paths, context=ctx)
        steps = [[db for db, _ in step] for step in case["steps"]]
        ouniq, oconn = kin_oracle.folding_graph(steps)
        assert list(uniq) == ouniq, case["name"]
        for t, step in enumerate(steps):
            n_prev = len(steps[t - 1])
            blk = conn[int(conn_off[t]):int(conn_off[t + 1])].reshape(len(step), n_prev)
            got = [list(np.nonzero(row)[0]) for row in blk]
            assert got == oconn[t], (case["name"], t)


def _ref_matrix(case):
    "the reference's transition matrix rebuilt from ITS connectivity lists (golden), rafft_kin.py:68-91"
    idx = {s: i for i, s in enumerate(case["struct_list"])}
    nrj = {}
    for step in case["steps"]:
        for db, e in step:
            nrj.setdefault(db, e)
    n = len(idx)
    M = np.zeros((n, n), dtype=np.longdouble)
    for t, step in enumerate(case["steps"]):
        prev = case["steps"][t - 1]
        for (db, _), lco in zip(step, case["conn"][t]):
            for si in lco:
                pdb = prev[si][0]
                d = nrj[db] - nrj[pdb]
                if idx[db] != idx[pdb]:
                    M[idx[pdb], idx[db]] = min(1.0, np.exp(-d / 0.61))
                    M[idx[db], idx[pdb]] = min(1.0, np.exp(d / 0.61))
    for i in range(n):
        M[i, i] = -M[i, :].sum()
    return M


@pytest.mark.gpu
def test_transition_matrix_bit_exact(ctx, kin_golden):
    "the part the GPU decides (graph -> rate matrix) is identical to the reference's, entry by entry"
    from rafft_b200 import rafft_kin as K
    for case in kin_golden:
        paths = _paths(case)
        flat = [st for step in paths for st in step]
        graph = K.folding_graph(paths, context=ctx)
        struct_list = [flat[i] for i in range(len(flat)) if graph[0][i] == i]
        smap = {st.str_struct: (si, st.energy) for si, st in enumerate(struct_list)}
        M = K.get_transition_mat(paths, len(struct_list), smap, graph)
        assert (M == _ref_matrix(case)).all(), case["name"]


@pytest.mark.gpu
def test_kinetics_populations_vs_reference(ctx, kin_golden):
    """Populations come out of scipy.linalg.eig/inv of that (identical) matrix on the host, as in the
    reference; they depend on the LAPACK build and CPU.  Tolerance: 1e-9 relative for the reference's
    example files (14 / 68 states; bit-identical in practice).  For the 874-state cfg-5 graph the
    reference's own formula W exp(Vt) W^-1 p is numerically unstable at late times (its golden final
    row already holds populations of -4e-3), so rows up to t = e^11 are compared at 1e-6 absolute
    and the later ones are not compared."""
    from rafft_b200 import rafft_kin as K
    for case in kin_golden:
        big = len(case["struct_list"]) > 200
        rtol, atol = (0.0, 1e-6) if big else (1e-9, 1e-12)
        traj, times, struct_list, equi = K.kinetics(_paths(case), case["max_time"], case["n_steps"])
        assert [s.str_struct for s in struct_list] == case["struct_list"]
        assert np.allclose(times, case["times"], rtol=1e-12)
        for r, want in zip(case["traj_rows"], case["traj"]):
            if big and r > case["n_steps"] // 2:
                continue
            assert np.allclose(np.asarray(traj[r], dtype=np.float64), want, rtol=rtol, atol=atol), (case["name"], r)
        for (s, e, p, i), (ws, we, wp, wi) in zip(equi, case["equi"]):
            assert s == ws and i == wi and e == we
            if not big:
                assert abs(float(p) - wp) <= rtol * abs(wp) + atol


@pytest.mark.gpu
def test_kin_edge_cases(ctx):
    from rafft_b200 import rafft_kin as K
    # one step, one structure: looks back at itself
    u, c, co, so = K.folding_graph([[_S("....", 0.0)]], context=ctx)
    assert list(u) == [0] and list(c) == [1]
    # pseudo-knot pile and duplicates across steps
    paths = [[_S("......", 0.0)], [_S("(....)", -1.0), _S("[....]", -1.0)], [_S("((..))", -2.0), _S("(....)", -1.0)]]
    u, c, co, so = K.folding_graph(paths, context=ctx)
    assert list(u) == [0, 1, 2, 3, 1]
    import kin_oracle
    _, oc = kin_oracle.folding_graph([[s.str_struct for s in st] for st in paths])
    flat = []
    for t, st in enumerate(paths):
        blk = c[int(co[t]):int(co[t + 1])].reshape(len(st), len(paths[t - 1]))
        flat.append([list(np.nonzero(r)[0]) for r in blk])
    assert flat == oc
    with pytest.raises(IndexError):
        K.folding_graph([[_S("..)", 0.0)]], context=ctx)


@pytest.mark.gpu
def test_kin_cli_matches_reference_output(kin_golden):
    "bin/rafft_kin example/rafft.out prints the reference's lines (populations to 3 decimals)"
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bin", "rafft_kin"), os.path.join(GOLD, "rafft.out")],
                         capture_output=True, text=True, check=True).stdout.strip().splitlines()
    case = kin_golden[0]
    want = sorted(case["equi"], key=lambda el: el[2])
    assert len(out) == len(want)
    for line, (s, e, p, i) in zip(out, want):
        assert line == "{} {:6.3f} {:5.1f} {:d}".format(s, p, e, i)
