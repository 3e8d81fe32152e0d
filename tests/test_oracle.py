"""The oracle (oracle/turner_eval.c, oracle/rafft_oracle.cpp) against every golden vector the
reference holds for the path.  CPU only."""
import os

import pytest

from conftest import GOLDEN, parse_rafft_text


def fmt(e_dcal):
    from rafft_b200.utils import energy_from_dcal
    return ("%6.1f" % energy_from_dcal(e_dcal)).strip()


# ---- energy evaluator ---------------------------------------------------------------------
def test_energy_kat_all_triples(oracle, energy_kat):
    "11,505 (sequence, structure, energy) triples of benchmark_results/*_scores.csv"
    assert len(energy_kat) == 11505
    bad = [(s, d, e, oracle.eval_db(s, d)) for s, d, e in energy_kat if oracle.eval_db(s, d) != e]
    assert not bad, bad[:3]


def test_vrna_eval_known_answer(oracle, kernel_kat):
    "rafft-rs/src/vienna.rs:92-109: pair table of the MFE structure -> -25.80 kcal/mol"
    k = kernel_kat["vrna_eval"]
    pt = k["pt_1based"]
    L = pt[0]
    db = "".join("." if pt[i] == 0 else "(" if pt[i] > i else ")" for i in range(1, L + 1))
    assert oracle.eval_db(k["seq"], db) == k["energy_dcal"] == -2580


@pytest.mark.parametrize("name", ["rafft.out", "rafft_20.out", "analysis_n50_ms5.txt"])
def test_energy_of_example_lines(oracle, name):
    seq, steps = parse_rafft_text(os.path.join(GOLDEN, name))
    for step in steps:
        for db, etxt in step:
            assert fmt(oracle.eval_db(seq, db)) == etxt.strip(), (db, etxt)


# ---- leaf kernels against the Rust unit-test constants --------------------------------------
def test_autocorrelation_kat(oracle, kernel_kat):
    k = kernel_kat["autocorrelation"]
    seq = k["seq"]
    got = oracle.auto_cor(seq, list(range(len(seq))), **k["weights"])
    assert len(got) == len(k["ac"]) == 2 * len(seq) - 1
    assert max(abs(a - b) for a, b in zip(got, k["ac"])) < 1e-6  # the Rust test's f32 tolerance


def test_consecutive_pairs_kat(oracle, kernel_kat):
    k = kernel_kat["consecutive_pairs"]
    seq = k["seq"]
    for lag, min_hp, nb, i, j, score in k["cases"]:
        assert oracle.window_slide(seq, list(range(len(seq))), lag, min_hp) == (nb, i, j, score), lag
    k = kernel_kat["consecutive_pairs_split"]
    for lag, min_hp, nb, i, j, score in k["cases"]:
        assert oracle.window_slide(k["seq"], k["pos"], lag, min_hp) == (nb, i, j, score), lag


# ---- end to end -------------------------------------------------------------------------
@pytest.mark.parametrize("name,kw", [("rafft.out", dict(nb_mode=100, max_stack=5)),
                                     ("rafft_20.out", dict(nb_mode=100, max_stack=20)),
                                     ("analysis_n50_ms5.txt", dict(nb_mode=50, max_stack=5))])
def test_example_trajectories(oracle, name, kw):
    "example/rafft.out, example/rafft_20.out, analysis.org:723-754: same text, same order"
    seq, steps = parse_rafft_text(os.path.join(GOLDEN, name))
    got = oracle.fold(seq, max_branch=1000, **kw)
    assert [[(db, fmt(e)) for db, e in st] for st in got] == [[(db, e.strip()) for db, e in st] for st in steps]


def test_against_unmodified_reference(oracle, ref_folds):
    "62 trajectories produced by the reference's own Python (tests/golden/ref_folds.json.gz)"
    assert len(ref_folds) >= 60
    for c in ref_folds:
        assert oracle.fold(c["seq"], **c["params"]) == c["steps"], c["name"]
