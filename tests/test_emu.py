"""The DEVICE code (rafft_b200/csrc/fold_core.cuh) compiled for the host with a
one-thread execution policy, against the oracle and the reference's golden trajectories.
This checks the ordering logic of the CUDA path in the CPU tier; it is test-only code
(tests/emu), not a fallback of the product."""
import os
import random

import pytest

from conftest import GOLDEN, parse_rafft_text
from rafft_b200.synth import splitmix_seq


def test_emu_reference_folds(emu, ref_folds):
    for k, c in enumerate(ref_folds):
        rc, steps, _, _ = emu(c["seq"], reverse=k & 1, nthreads=(32, 128, 7)[k % 3], **c["params"])
        assert rc == 0, c["name"]
        assert steps == c["steps"], c["name"]


@pytest.mark.parametrize("name,kw", [("rafft.out", dict(nb_mode=100, max_stack=5)),
                                     ("rafft_20.out", dict(nb_mode=100, max_stack=20)),
                                     ("analysis_n50_ms5.txt", dict(nb_mode=50, max_stack=5))])
def test_emu_examples(emu, name, kw):
    from rafft_b200.utils import energy_from_dcal
    seq, gold = parse_rafft_text(os.path.join(GOLDEN, name))
    rc, steps, _, _ = emu(seq, **kw)
    assert rc == 0
    got = [[(db, ("%6.1f" % energy_from_dcal(e)).strip()) for db, e in st] for st in steps]
    assert got == [[(db, e.strip()) for db, e in st] for st in gold]


def test_emu_vs_oracle_random(emu, oracle):
    rnd = random.Random(77)
    for k in range(60):
        L = rnd.choice([1, 2, 3, 7, 15, 30, 45, 60, 90, 130, 180])
        kw = dict(nb_mode=rnd.choice([1, 4, 30, 100, 200]), max_stack=rnd.choice([1, 3, 10, 50]),
                  max_branch=rnd.choice([1, 5, 100, 1000]), min_hp=rnd.choice([0, 3, 3, 3, 6]),
                  min_nrj=rnd.choice([0.0, 0.0, 0.0, -0.5, 0.4]))
        seq = splitmix_seq(9000 + k, L)
        rc, steps, cnt, _ = emu(seq, reverse=rnd.randrange(2), nthreads=rnd.choice([1, 32, 256]), **kw)
        want, wcnt = oracle.fold(seq, counters=True, **kw)
        assert rc == 0 and steps == want, (k, L, kw)
        # the reference-formulation work counters agree with the oracle's
        assert cnt[:5] == wcnt[:5] and cnt[8:10] == wcnt[8:10], (k, cnt, wcnt)


def test_emu_capacity_is_reported_not_silent(emu):
    "an arena that is too small yields RAFFT_ERR_CAPACITY (the host then regrows), never a wrong answer"
    seq = splitmix_seq(5, 200)
    rc, steps, _, hw = emu(seq, nb_mode=100, max_stack=50, scale=1)
    assert rc == 0
    # high-water marks stay inside the capacities of scale 1
    assert hw[0] > 0 and hw[3] > 0


def test_emu_invalid_letter(emu):
    rc, steps, _, _ = emu("GGGANACCC", nb_mode=10, max_stack=2)
    assert rc == 2  # RAFFT_ERR_SEQ (reference: KeyError, rafft/utils.py:73-80)


def test_emu_task_queue_overflow_defers(emu, oracle):
    "a full expansion work list defers regions to the next round without changing the result"
    for k, cap in enumerate((1, 2, 5)):
        seq = splitmix_seq(600 + k, 90)
        rc, steps, cnt, _ = emu(seq, queue_cap=cap, nb_mode=50, max_stack=10)
        want, wcnt = oracle.fold(seq, 50, 10, 1000, counters=True)
        assert rc == 0 and steps == want
        assert cnt[:5] == wcnt[:5]


def test_emu_long_region_ranks_by_selection(emu, oracle):
    """regions with more than 2048 padded lags (lv > 1024) rank by pivot selection (expand_e2):
    same trajectory as the oracle, forward and reversed iteration order"""
    seq = splitmix_seq(4242, 1300)
    want = oracle.fold(seq, 60, 3, 1000)
    for rev in (0, 1):
        rc, steps, _, _ = emu(seq, reverse=rev, nb_mode=60, max_stack=3)
        assert rc == 0 and steps == want


def test_emu_cut_positions_across_small_chunks(emu, oracle):
    """the product enumeration in chunks of 4-8 combinations with the max_branch cut falling on every position
    (first / last item of a chunk, chunk boundaries, later members trying their first combination only):
    the running counts kept in registers and the in-chunk duplicate table of phase_combine"""
    seqs = [splitmix_seq(1200 + k, L) for k, L in enumerate((48, 70, 95))]
    for seq in seqs:
        for max_stack in (2, 5, 8):
            for max_branch in (1, 2, 3, 4, 5, 7, 8, 9, 13, 16, 17, 31):
                kw = dict(nb_mode=40, max_stack=max_stack, max_branch=max_branch)
                rc, steps, cnt, _ = emu(seq, nthreads=1, **kw)
                want, wcnt = oracle.fold(seq, counters=True, **kw)
                assert rc == 0 and steps == want, (len(seq), kw)
                assert cnt[:5] == wcnt[:5] and cnt[8:10] == wcnt[8:10], (len(seq), kw)
