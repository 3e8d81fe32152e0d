"""The Python surface above the C-ABI without a GPU: the lazy per-sequence view of a result
(BatchFolds), the oracle's full-batch comparator, the reference's import names (`rafft`), and the
ranges the single-process multi-device entry deals out.  A result in the layout of
include/rafft_b200.h is packed from ORACLE folds here, so no compute call reaches the library."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class PackedResult:
    "the arrays of a rafft_result, filled from oracle step lists (test stand-in for RawResult)"

    def __init__(self, seqs, steps_per_seq, traj=True):
        n = len(seqs)
        self.n_seq = n
        self.status = np.zeros(n, dtype=np.int32)
        self.lengths = np.array([len(s) for s in seqs], dtype=np.int64)
        so, do, to, eo = [0], [0], [0], [0]
        energy, db, entries = [], [], []
        for s, steps in zip(seqs, steps_per_seq):
            index = {}
            for step in (steps if traj else steps[-1:]):
                for d, e in step:
                    if d not in index:
                        index[d] = len(index)
                        energy.append(e)
                        db.append(d)
                    entries.append(index[d])
                eo.append(len(entries))
            so.append(len(energy))
            do.append(do[-1] + len(index) * len(s))
            to.append(len(eo) - 1)
        self.struct_off, self.db_off = np.array(so, dtype=np.int64), np.array(do, dtype=np.int64)
        self.step_off, self.entry_off = np.array(to, dtype=np.int64), np.array(eo, dtype=np.int64)
        self.energy = np.array(energy, dtype=np.int32)
        self.db = np.frombuffer("".join(db).encode(), dtype=np.uint8).copy()
        self.entry_struct = np.array(entries, dtype=np.int32)
        self.n_struct = len(energy)


@pytest.fixture(scope="module")
def batch(oracle):
    from rafft_b200.synth import config_batch
    seqs = config_batch("cfg2", 12) + ["G", "GGGAAACCC"]
    steps, _ = oracle.fold_batch(seqs, 50, 6, 100)
    return seqs, steps


def test_batchfolds_is_lazy_view(batch):
    from rafft_b200.rafft import BatchFolds
    seqs, steps = batch
    for traj in (True, False):
        bf = BatchFolds(PackedResult(seqs, steps, traj), traj)
        assert len(bf) == len(seqs)
        for i, r in enumerate(bf):
            final, tr = r if traj else (r, None)
            assert [(s.str_struct, s.energy_dcal) for s in final] == steps[i][-1]
            if traj:
                assert [[(s.str_struct, s.energy_dcal) for s in st] for st in tr] == steps[i]
        last = bf[-1][0] if traj else bf[-1]
        assert [(s.str_struct, s.energy_dcal) for s in last] == steps[-1][-1]
        dbs, es = bf.best()
        assert dbs == [st[-1][0][0] for st in steps] and es.tolist() == [st[-1][0][1] for st in steps]
        with pytest.raises(IndexError):
            bf[len(seqs)]
    # energies print like the reference: %6.1f of the float32-valued energy
    bf = BatchFolds(PackedResult(seqs, steps, False), False)
    assert all(abs(s.energy - s.energy_dcal / 100.0) < 1e-5 for s in bf[0])


def test_failed_sequence_raises(batch):
    from rafft_b200 import _native as N
    from rafft_b200.rafft import BatchFolds
    seqs, steps = batch
    pr = PackedResult(seqs, steps)
    pr.status[3] = N.RAFFT_ERR_CAPACITY
    with pytest.raises(N.NativeError):
        BatchFolds(pr, True)


def test_oracle_batch_check_detects_differences(batch, oracle):
    "the comparator the full-batch GPU parity tests rely on: silent on equal results, exact on which differ"
    seqs, steps = batch
    for traj in (True, False):
        pr = PackedResult(seqs, steps, traj)
        bad, cnt = oracle.fold_batch_check(seqs, pr, 50, 6, 100, traj=traj)
        assert bad == [] and cnt.shape == (len(seqs), cnt.shape[1])
        pr.energy[int(pr.struct_off[5])] += 1                    # one energy of sequence 5
        k = int(pr.db_off[9]) + 3
        pr.db[k] = ord("(") if pr.db[k] == ord(".") else ord(".")   # one character of sequence 9
        assert oracle.fold_batch_check(seqs, pr, 50, 6, 100, traj=traj)[0] == [5, 9]
    pr = PackedResult(seqs, steps, True)
    e0, e1 = int(pr.entry_off[int(pr.step_off[2]) + 1]), int(pr.entry_off[int(pr.step_off[2]) + 2])
    if e1 - e0 >= 2:                                              # beam order of one step of sequence 2
        pr.entry_struct[e0], pr.entry_struct[e0 + 1] = pr.entry_struct[e0 + 1], pr.entry_struct[e0]
        assert oracle.fold_batch_check(seqs, pr, 50, 6, 100)[0] == [2]
    assert oracle.fold_batch_check(seqs, PackedResult(seqs, steps), 50, 7, 100)[0] != []  # other parameters


def test_reference_import_names():
    "README.org:171-180 style user code imports unchanged (rafft/__init__.py:1-2 of the reference)"
    import rafft
    from rafft import fold, kinetics
    from rafft.rafft import fold as fold2
    from rafft.rafft_kin import kinetics as kin2
    from rafft.utils import Structure, dot_bracket, paired_positions, parse_rafft_output, read_fasta  # noqa: F401
    import rafft_b200
    assert fold is fold2 is rafft_b200.fold and kinetics is kin2 and rafft.fold is fold


def test_letters_checked_in_one_pass():
    from rafft_b200.rafft import _check_letters
    _check_letters(b"ACGUUGCA" * 1000)
    _check_letters(b"")
    for bad in (b"ACGN", b"acgu", b"ACGT", b"AC GU"):
        with pytest.raises(KeyError):
            _check_letters(bad)


def test_multi_device_ranges(native_lib_path):
    "rafft_fold_batch_multi deals contiguous, non-empty, cost-balanced ranges of the input order"
    from rafft_b200.shard import estimated_cost, plan_ranges
    from rafft_b200.synth import config_batch
    lens = [len(s) for s in config_batch("cfg2", 3000)]
    cost = np.array([estimated_cost(l, 200) for l in lens])
    for parts in (1, 2, 3, 4, 8):
        lo = plan_ranges(lens, parts, 200)
        assert lo[0] == 0 and lo[-1] == len(lens) and (np.diff(lo) > 0).all()
        share = np.array([cost[a:b].sum() for a, b in zip(lo[:-1], lo[1:])])
        assert share.max() / share.mean() < 1.01
    assert plan_ranges([100, 100, 100], 3).tolist() == [0, 1, 2, 3]
    assert plan_ranges([500, 10, 10, 10], 4).tolist() == [0, 1, 2, 3, 4]
    assert plan_ranges([10, 10, 10, 500], 2).tolist() == [0, 3, 4]
    assert plan_ranges([], 2).tolist() == [0, 0, 0]


def test_parameter_file_roundtrip(tmp_path, native_lib_path):
    """ViennaRNA `## RNAfold parameter file v2.0` reader (the reference loads such files,
    rafft-rs/src/vienna.rs:75-81): the built-in tables written out and read back are identical;
    comments, INF and a changed entry survive; malformed files are refused"""
    from rafft_b200 import _native as N
    from rafft_b200 import params as P
    ep = N.default_energy_params()
    path = str(tmp_path / "t.par")
    P.write_par(ep, path)
    ep2 = P.read_par(path)
    assert bytes(ep) == bytes(ep2)
    txt = open(path).read()
    assert "# int22" in txt and "INF" in txt and txt.startswith("## RNAfold parameter file v2.0")
    # ViennaRNA's own files carry /* */ comments between and inside the rows
    txt = txt.replace("# stack\n", "# stack\n/* a comment\n over two lines */\n").replace("  -240  -330", "  -250 /* CG CG */  -330", 1)
    open(path, "w").write(txt)
    ep3 = P.read_par(path)
    assert ep3.stack[0] == -250 and ep3.stack[1] == -330 and list(ep3.int22) == list(ep.int22)
    assert ep3.hairpin[0] == P.INF and ep3.terminal_au == ep.terminal_au and ep3.lxc == ep.lxc
    assert [ep3.tetra[k].value for k in range(ep3.n_tetra)] == [ep.tetra[k].value for k in range(ep.n_tetra)]
    open(path, "w").write(txt.replace("# int21", "# int21_renamed"))
    with pytest.raises(ValueError):
        P.read_par(path)
    open(path, "w").write("not a parameter file\n")
    with pytest.raises(ValueError):
        P.read_par(path)
