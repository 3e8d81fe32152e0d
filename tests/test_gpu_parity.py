"""Parity of the CUDA path (through the C-ABI, rafft_b200/librafft_b200.so) with the oracle
and with the reference's golden vectors.  Bar: bit-exact -- same dot-brackets, same integer
energies (dcal/mol), same trajectory order."""
import os
import random
import subprocess
import sys

import numpy as np
import pytest

from conftest import GOLDEN, ROOT, parse_rafft_text
from rafft_b200.synth import config_batch, splitmix_seq

pytestmark = pytest.mark.gpu


def gpu_steps(ctx, seqs, **kw):
    from rafft_b200 import rafft as R
    p = dict(nb_mode=100, max_stack=1, max_branch=1000, min_hp=3, min_nrj=0.0, gc=3.0, au=2.0, gu=1.0)
    p.update(kw)
    offs = np.zeros(len(seqs) + 1, dtype=np.int64)
    np.cumsum([len(s) for s in seqs], out=offs[1:])
    prm = R.make_params(p["nb_mode"], p["max_stack"], p["max_branch"], p["min_hp"], p["min_nrj"], True, 37.0,
                        p["gc"], p["au"], p["gu"])
    raw = ctx.fold_raw("".join(seqs).encode(), offs, prm)
    assert (raw.status == 0).all(), raw.status
    return [raw.steps(i) for i in range(len(seqs))], raw


def test_native_library_is_loaded(ctx):
    "the GPU tier runs the in-tree CUDA library, not a fallback"
    maps = open("/proc/self/maps").read()
    assert "librafft_b200.so" in maps


# ---- (d) energy evaluator --------------------------------------------------------------
def test_eval_structures_energy_kat(ctx, energy_kat, kernel_kat):
    seqs = [s for s, _, _ in energy_kat]
    dbs = [d for _, d, _ in energy_kat]
    got = ctx.eval_structures(seqs, dbs)
    want = np.array([e for _, _, e in energy_kat], dtype=np.int32)
    assert (got == want).all(), np.nonzero(got != want)[0][:5]
    k = kernel_kat["vrna_eval"]
    pt = k["pt_1based"]
    db = "".join("." if pt[i] == 0 else "(" if pt[i] > i else ")" for i in range(1, pt[0] + 1))
    assert ctx.eval_structures([k["seq"]], [db])[0] == -2580


def test_eval_structures_vs_oracle_random(ctx, oracle):
    "random structures taken from oracle trajectories + malformed input"
    seqs, dbs = [], []
    for k in range(40):
        s = splitmix_seq(400 + k, 40 + 5 * k)
        for st in oracle.fold(s, 50, 5, 100)[-1]:
            seqs.append(s)
            dbs.append(st[0])
    got = ctx.eval_structures(seqs, dbs)
    assert [int(g) for g in got] == [oracle.eval_db(s, d) for s, d in zip(seqs, dbs)]
    bad = ctx.eval_structures(["GGGAAACCC", "GGGAAACCC", "GGGAAACCC"], ["(((...)).", "((.....))", "(((...)))"])
    assert bad[0] == -(2 ** 31) and bad[1] == oracle.eval_db("GGGAAACCC", "((.....))")


# ---- (a)-(c),(e) whole path against golden files ------------------------------------------
@pytest.mark.parametrize("name,kw", [("rafft.out", dict(nb_mode=100, max_stack=5)),
                                     ("rafft_20.out", dict(nb_mode=100, max_stack=20)),
                                     ("analysis_n50_ms5.txt", dict(nb_mode=50, max_stack=5))])
def test_example_files(ctx, name, kw):
    from rafft_b200.utils import energy_from_dcal
    seq, gold = parse_rafft_text(os.path.join(GOLDEN, name))
    steps = gpu_steps(ctx, [seq], **kw)[0][0]
    got = [[(db, ("%6.1f" % energy_from_dcal(e)).strip()) for db, e in st] for st in steps]
    assert got == [[(db, e.strip()) for db, e in st] for st in gold]


def test_reference_python_trajectories(ctx, ref_folds):
    "62 trajectories of the unmodified reference Python, incl. edge cases (L=1, polyA, mb=1, ...)"
    by_params = {}
    for c in ref_folds:
        by_params.setdefault(tuple(sorted(c["params"].items())), []).append(c)
    for key, cases in by_params.items():
        got, _ = gpu_steps(ctx, [c["seq"] for c in cases], **dict(key))
        for c, g in zip(cases, got):
            assert g == c["steps"], c["name"]


def test_kernel_kats_through_the_path(ctx, kernel_kat, oracle):
    """The Rust stem-scan/correlation KAT sequences: the first folding step of the CUDA path
    must pick exactly the stems the oracle's auto_cor + window_slide produce."""
    for seq in (kernel_kat["consecutive_pairs"]["seq"], kernel_kat["autocorrelation"]["seq"],
                kernel_kat["consecutive_pairs_split"]["seq"]):
        for n in (1, 3, 10, 100):
            assert gpu_steps(ctx, [seq], nb_mode=n, max_stack=30)[0][0] == oracle.fold(seq, n, 30, 1000)


# ---- batches vs oracle -----------------------------------------------------------------
def fold_and_check(ctx, oracle, seqs, nb_mode, max_stack, traj=True, counters=True):
    """the whole batch through the C-ABI, then every beam of every sequence (energies, dot-brackets,
    order; with traj every folding step) against the oracle's own fold, compared inside the oracle
    library; the reference-formulation work counters too"""
    from rafft_b200 import rafft as R
    offs = np.zeros(len(seqs) + 1, dtype=np.int64)
    np.cumsum([len(s) for s in seqs], out=offs[1:])
    raw = ctx.fold_raw("".join(seqs).encode(), offs, R.make_params(nb_mode, max_stack, 1000, traj=traj))
    assert (raw.status == 0).all(), np.bincount(raw.status)
    bad, cnt = oracle.fold_batch_check(seqs, raw, nb_mode, max_stack, 1000, traj=traj)
    assert not bad, "%d sequences differ from the oracle, first %s" % (len(bad), bad[:5])
    if counters:
        assert (raw.counters[:, :5] == cnt[:, :5]).all() and (raw.counters[:, 8:10] == cnt[:, 8:10]).all()
    return raw


def test_full_batch_cfg2_vs_oracle(ctx, oracle):
    "BASELINE configs[1]: ALL 10,000 sequences at -n 100 -ms 50, whole trajectories, bit-exact"
    raw = fold_and_check(ctx, oracle, config_batch("cfg2"), 100, 50)
    assert raw.n_seq == 10000


def test_batch_cfg3_2500_vs_oracle(ctx, oracle):
    "BASELINE configs[2]: every 4th sequence of the 10k batch at -n 200 -ms 200, whole trajectories"
    fold_and_check(ctx, oracle, config_batch("cfg2")[1::4], 200, 200)


def test_cfg3_full_batch_final_beams_vs_oracle(ctx, oracle):
    "BASELINE configs[2] as benchmarked (no trajectory): the final 200 structures of a 1,250-sequence GPU share"
    fold_and_check(ctx, oracle, config_batch("cfg2")[3::8], 200, 200, traj=False)


def test_lanes_and_loop_modes_give_identical_results(ctx):
    """one lane / two lanes (two priority streams, result copy of the first overlapping the fold of the
    second), graph WHILE loop / stream loop with look-ahead / stepwise profile mode, fused beam kernel /
    separate select, materialise and gather kernels: same bytes"""
    from rafft_b200 import rafft as R
    seqs = config_batch("cfg2", 1500)
    offs = np.zeros(len(seqs) + 1, dtype=np.int64)
    np.cumsum([len(s) for s in seqs], out=offs[1:])
    sb, prm = "".join(seqs).encode(), R.make_params(100, 50, 1000, traj=True)
    ref = None
    try:
        for lanes, loop, prof, fuse in ((1, 0, 0, 1), (2, 0, 0, 1), (2, 1, 0, 0), (1, 1, 0, 1), (1, 0, 1, 1), (1, 0, 1, 0),
                                        (1, 0, 0, 0), (2, 0, 0, 1)):
            ctx.set_option("lanes", lanes); ctx.set_option("loop", loop); ctx.set_option("profile", prof)
            ctx.set_option("fuse", fuse)
            raw = ctx.fold_raw(sb, offs, prm)
            assert raw.lanes == lanes and raw.loop_mode == (2 if prof else loop)
            cur = [np.array(a) for a in (raw.status, raw.struct_off, raw.energy, raw.db_off, raw.db, raw.step_off,
                                         raw.entry_off, raw.entry_struct)]
            if prof:
                assert raw.family_ms["expand_fast"] > 0 and raw.family_launches["combine"] == raw.iterations
            if ref is None:
                ref = cur
            else:
                assert all((a == b).all() for a, b in zip(ref, cur)), (lanes, loop, prof, fuse)
            raw.close()
    finally:
        ctx.set_option("lanes", 0); ctx.set_option("loop", 0); ctx.set_option("profile", 0); ctx.set_option("fuse", 0)


def test_multi_device_entry_matches_single(ctx, oracle):
    """rafft_fold_batch_multi (one process, one host thread + stream per device): identical arrays to the
    single-device call on every GPU count the box has; with one GPU the entry is exercised with one context"""
    import torch
    from rafft_b200 import rafft as R
    seqs = config_batch("cfg2", 900)
    offs = np.zeros(len(seqs) + 1, dtype=np.int64)
    np.cumsum([len(s) for s in seqs], out=offs[1:])
    sb, prm = "".join(seqs).encode(), R.make_params(100, 20, 1000, traj=True)
    one = ctx.fold_raw(sb, offs, prm)
    ref = [np.array(a) for a in (one.status, one.struct_off, one.energy, one.db_off, one.db, one.step_off,
                                 one.entry_off, one.entry_struct, one.counters[:, :10])]
    ndev = torch.cuda.device_count()
    for nd in sorted({1, min(2, ndev), ndev}):
        raw = R.fold_raw_multi([R.get_context(d) for d in range(nd)], sb, offs, prm)
        cur = [np.array(a) for a in (raw.status, raw.struct_off, raw.energy, raw.db_off, raw.db, raw.step_off,
                                     raw.entry_off, raw.entry_struct, raw.counters[:, :10])]
        assert all(a.shape == b.shape and (a == b).all() for a, b in zip(ref, cur)), nd
        raw.close()
    got = R.fold_batch(seqs[:40], 100, 20, 1000, traj=True, devices=list(range(ndev)))
    want, _ = oracle.fold_batch(seqs[:40], 100, 20, 1000)
    assert [[[(s.str_struct, s.energy_dcal) for s in st] for st in tr] for _, tr in got] == want


def test_random_parameters_vs_oracle(ctx, oracle):
    rnd = random.Random(4242)
    for k in range(25):
        kw = dict(nb_mode=rnd.choice([1, 5, 50, 100, 200]), max_stack=rnd.choice([1, 2, 10, 50]),
                  max_branch=rnd.choice([1, 7, 100, 1000]), min_hp=rnd.choice([0, 3, 3, 5]),
                  min_nrj=rnd.choice([0.0, 0.0, -0.6, 0.35]),
                  **rnd.choice([dict(), dict(gc=1.0, au=1.0, gu=1.0), dict(gc=2.5, au=1.5, gu=0.5)]))
        seqs = [splitmix_seq(7000 + 31 * k + j, rnd.choice([1, 2, 9, 33, 70, 121, 200])) for j in range(12)]
        got, _ = gpu_steps(ctx, seqs, **kw)
        want, _ = oracle.fold_batch(seqs, **kw)
        assert got == want, kw


def test_edge_cases(ctx, oracle):
    from rafft_b200 import _native as N
    from rafft_b200 import rafft as R
    seqs = ["G", "GC", "A" * 50, "GC" * 40, "G" * 45 + "AAAA" + "C" * 45, "GGGAAACCC"]
    got, _ = gpu_steps(ctx, seqs, nb_mode=100, max_stack=10)
    assert got == oracle.fold_batch(seqs, 100, 10, 1000)[0]
    # empty batch
    raw = ctx.fold_raw(b"", np.zeros(1, dtype=np.int64), R.make_params(traj=True))
    assert raw.n_seq == 0 and raw.n_struct == 0
    # a bad letter fails that sequence only (reference: KeyError), neighbours are unaffected
    offs = np.array([0, 9, 18, 27], dtype=np.int64)
    raw = ctx.fold_raw(b"GGGAAACCCGGGANACCCGGGAAACCC", offs, R.make_params(max_stack=3, traj=True))
    assert list(raw.status) == [0, N.RAFFT_ERR_SEQ, 0] and raw.steps(0) == raw.steps(2) != []
    # unsupported arguments are refused, not approximated
    for bad in (dict(temp=25.0), dict(gc_wei=0.3), dict(max_stack=0)):
        with pytest.raises(N.NativeError):
            ctx.fold_raw(b"GGGAAACCC", np.array([0, 9], dtype=np.int64), R.make_params(**bad))


def test_traj_off_returns_final_beam(ctx):
    from rafft_b200 import fold_batch
    seqs = config_batch("cfg2", 40)
    a = fold_batch(seqs, 100, 20, 1000, traj=False, context=ctx)
    b = fold_batch(seqs, 100, 20, 1000, traj=True, context=ctx)
    for x, (fin, tr) in zip(a, b):
        assert [(s.str_struct, s.energy_dcal) for s in x] == [(s.str_struct, s.energy_dcal) for s in fin]
        assert [(s.str_struct, s.energy_dcal) for s in tr[-1]] == [(s.str_struct, s.energy_dcal) for s in fin]


def test_long_sequences_vs_oracle(ctx, oracle):
    "1000-nt and a 3000-nt (config 4 scale) sequence at reduced -ms so the oracle ends in seconds"
    seqs = [splitmix_seq(3000, 3000), splitmix_seq(1000, 1000)]
    got, _ = gpu_steps(ctx, seqs, nb_mode=200, max_stack=10)
    assert got == oracle.fold_batch(seqs, 200, 10, 1000)[0]


# ---- size-independent properties at full size ------------------------------------------------
def test_full_batch_properties(ctx):
    "all 10k sequences of config 2: structural invariants + energies re-derived by the evaluator kernel"
    seqs = config_batch("cfg2")
    from rafft_b200 import rafft as R
    offs = np.zeros(len(seqs) + 1, dtype=np.int64)
    np.cumsum([len(s) for s in seqs], out=offs[1:])
    raw = ctx.fold_raw("".join(seqs).encode(), offs, R.make_params(100, 50, 1000, traj=False))
    assert (raw.status == 0).all()
    ns = np.diff(raw.struct_off)
    assert (ns >= 1).all() and (ns <= 50).all()
    ev_s, ev_d, ev_e = [], [], []
    for i in range(0, len(seqs), 7):
        final = raw.steps(i)[-1]
        es = [e for _, e in final]
        assert es == sorted(es)                       # beam sorted by energy
        assert len(set(db for db, _ in final)) == len(final)  # no duplicates (global seen set)
        for db, e in final:
            assert len(db) == len(seqs[i]) and db.count("(") == db.count(")")
            ev_s.append(seqs[i]); ev_d.append(db); ev_e.append(e)
    got = ctx.eval_structures(ev_s, ev_d)               # incremental dE sums == full evaluation
    assert (got == np.array(ev_e, dtype=np.int32)).all()
    # determinism / idempotence
    raw2 = ctx.fold_raw("".join(seqs).encode(), offs, R.make_params(100, 50, 1000, traj=False))
    assert (raw2.db == raw.db).all() and (raw2.energy == raw.energy).all()


# ---- the drop-in surface ---------------------------------------------------------------
def test_cli_reproduces_example_file():
    seq = open(os.path.join(GOLDEN, "rafft.out")).readline().strip()
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bin", "rafft"), "-s", seq, "-ms", "5", "--traj"],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    assert out.stdout == open(os.path.join(GOLDEN, "rafft.out")).read()


def test_cli_batch_fasta_and_bench_line(tmp_path, oracle):
    """wire formats: a multi-record FASTA folded as one batch; the --bench line of
    bin/rafft:66 (`seq len struct E nbp`) and the final-structure blocks, against the oracle"""
    from rafft_b200.synth import splitmix_seq
    from rafft_b200.utils import energy_from_dcal
    seqs = [splitmix_seq(700 + k, L) for k, L in enumerate((31, 64, 97, 150))]
    fa = tmp_path / "in.fa"
    fa.write_text("".join(">s%d some text\n%s\n%s\n" % (k, s[:40].replace("U", "T"), s[40:]) for k, s in enumerate(seqs)))
    run = lambda *a: subprocess.run([sys.executable, os.path.join(ROOT, "bin", "rafft"), "-sf", str(fa), "--batch",
                                     "-n", "50", "-ms", "4"] + list(a), capture_output=True, text=True, timeout=300)
    want = [oracle.fold(s, 50, 4, 1000)[-1] for s in seqs]          # last beam: [(db, dcal)]
    out = run("--bench")
    assert out.returncode == 0, out.stderr
    lines = ["%s %d %s %6.1f %d" % (s, len(s), db, energy_from_dcal(e), db.count("(")) for s, beam in zip(seqs, want) for db, e in beam]
    assert out.stdout.splitlines() == lines
    out = run()
    assert out.returncode == 0, out.stderr
    lines = [l for s, beam in zip(seqs, want) for l in [s] + ["%s %6.1f" % (db, energy_from_dcal(e)) for db, e in beam]]
    assert out.stdout.splitlines() == lines


def test_fold_api_like_reference(ctx):
    from rafft_b200 import fold
    seq = open(os.path.join(GOLDEN, "rafft_20.out")).readline().strip()
    final, traj = fold(seq, max_stack=20, traj=True, max_branch=1000)
    _, gold = parse_rafft_text(os.path.join(GOLDEN, "rafft_20.out"))
    assert [s.str_struct for s in final] == [db for db, _ in gold[-1]]
    assert ["%6.1f" % s.energy for s in traj[1]] == ["%6.1f" % float(e) for _, e in gold[1]]
    assert isinstance(fold(seq), list) and len(fold(seq)) == 1


# ---- the BASELINE.json configurations at their own parameters ------------------------------
def test_cfg1_single_120nt(ctx, oracle):
    "configs[0]: one synthetic 120-nt RNA, -n 100 -ms 50 --traj, against the CPU run of the oracle"
    s = splitmix_seq(1, 120)
    got, _ = gpu_steps(ctx, [s], nb_mode=100, max_stack=50)
    assert got[0] == oracle.fold(s, 100, 50, 1000)


def test_cfg3_full_batch_properties(ctx):
    "configs[2] at full size (10k sequences, -n 200 -ms 200): invariants + energies re-derived"
    from rafft_b200 import rafft as R
    seqs = config_batch("cfg2")
    offs = np.zeros(len(seqs) + 1, dtype=np.int64)
    np.cumsum([len(s) for s in seqs], out=offs[1:])
    raw = ctx.fold_raw("".join(seqs).encode(), offs, R.make_params(200, 200, 1000, traj=False))
    assert (raw.status == 0).all()
    ns = np.diff(raw.struct_off)
    assert (ns >= 1).all() and (ns <= 200).all()
    ev_s, ev_d, ev_e = [], [], []
    for i in range(0, len(seqs), 41):
        final = raw.steps(i)[-1]
        es = [e for _, e in final]
        assert es == sorted(es) and len(set(db for db, _ in final)) == len(final)
        for db, e in final[::5]:
            assert len(db) == len(seqs[i]) and db.count("(") == db.count(")")
            ev_s.append(seqs[i]); ev_d.append(db); ev_e.append(e)
    got = ctx.eval_structures(ev_s, ev_d)
    assert (got == np.array(ev_e, dtype=np.int32)).all()


def test_cfg4_3000nt_vs_oracle(ctx, oracle):
    "configs[3] parameters (-n 200 -ms 200) on eight 3,000-nt inputs: whole trajectories bit-exact against the oracle"
    fold_and_check(ctx, oracle, [splitmix_seq(3000 + k, 3000) for k in range(8)], 200, 200)


def test_cfg5_fold_then_kinetics(ctx, oracle):
    "configs[4]: 200-nt RNA at -ms 200 --traj, the GPU-built trajectory fed to the kinetics"
    from rafft_b200 import rafft as R, rafft_kin as K
    s = splitmix_seq(200, 200)
    final, traj = R.fold_batch([s], 100, 200, 1000, traj=True, context=ctx)[0]
    want = oracle.fold(s, 100, 200, 1000)
    assert [[(st.str_struct, st.energy_dcal) for st in step] for step in traj] == want
    pops, times, struct_list, equi = K.kinetics(traj, 30, 100)
    assert len(times) == 101 and len(struct_list) == len({st.str_struct for step in traj for st in step})
    early = np.asarray(pops[50], dtype=np.float64)
    assert abs(early.sum() - 1.0) < 1e-9 and early.min() > -1e-6


def test_seen_generation_wrap(oracle):
    """The `seen` table is generation-tagged (no clean-up between folds of an arena): folds that
    reuse the arenas, across the wrap of the 16-bit generation counter, stay bit-exact."""
    code = r'''
import sys, numpy as np
sys.path.insert(0, %r); sys.path.insert(0, %r)
import pyoracle
from rafft_b200 import rafft as R
from rafft_b200.synth import splitmix_seq
ctx = R.FoldContext(0)
seqs = [splitmix_seq(900 + k, 40 + 7 * k) for k in range(24)]
want, _ = pyoracle.fold_batch(seqs, 100, 20, 1000)
offs = np.zeros(len(seqs) + 1, dtype=np.int64); np.cumsum([len(s) for s in seqs], out=offs[1:])
prm = R.make_params(100, 20, 1000, traj=True)
for rep in range(5):                      # generations 65534, 65535, wrap -> 1, 2, 3
    order = seqs if rep %% 2 == 0 else seqs[::-1]
    o = np.zeros(len(order) + 1, dtype=np.int64); np.cumsum([len(s) for s in order], out=o[1:])
    raw = ctx.fold_raw("".join(order).encode(), o, prm)
    assert (raw.status == 0).all()
    w = want if rep %% 2 == 0 else want[::-1]
    for i in range(len(order)):
        assert raw.steps(i) == w[i], (rep, i)
    raw.close()
ctx.close()
print("ok")
''' % (ROOT, os.path.join(ROOT, "oracle"))
    env = dict(os.environ, RAFFT_SEEN_GEN0="65533")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env)
    assert out.returncode == 0 and out.stdout.strip() == "ok", out.stderr[-2000:]


def test_parameter_file_context(ctx, tmp_path):
    "a context built from a ViennaRNA parameter file: the built-in tables written out fold identically; a changed stack entry changes energies"
    from rafft_b200 import _native as N
    from rafft_b200 import params as P
    from rafft_b200 import rafft as R
    path = str(tmp_path / "t.par")
    ep = N.default_energy_params()
    P.write_par(ep, path)
    seqs = config_batch("cfg2", 30)
    want, _ = gpu_steps(ctx, seqs, nb_mode=100, max_stack=10)
    c2 = R.FoldContext(0, P.read_par(path))
    got, _ = gpu_steps(c2, seqs, nb_mode=100, max_stack=10)
    assert got == want
    c2.close()
    ep.stack[1] -= 100                      # CG followed by GC, 1 kcal/mol more stable
    ep.stack[6] -= 100                      # and its mirror entry
    P.write_par(ep, path)
    c3 = R.FoldContext(0, P.read_par(path))
    other, _ = gpu_steps(c3, seqs, nb_mode=100, max_stack=10)
    assert other != want
    c3.close()


def test_cli_devices_flag(tmp_path, oracle):
    "bin/rafft --devices all: one process folds the FASTA records on every GPU of the box, output in input order"
    from rafft_b200.utils import energy_from_dcal
    seqs = [splitmix_seq(800 + k, L) for k, L in enumerate((60, 45, 120, 33, 90, 75))]
    fa = tmp_path / "in.fa"
    fa.write_text("".join(">s%d\n%s\n" % (k, s) for k, s in enumerate(seqs)))
    want = [oracle.fold(s, 50, 3, 1000)[-1] for s in seqs]
    lines = [l for s, beam in zip(seqs, want) for l in [s] + ["%s %6.1f" % (db, energy_from_dcal(e)) for db, e in beam]]
    for dev in ("all", "0"):
        out = subprocess.run([sys.executable, os.path.join(ROOT, "bin", "rafft"), "-sf", str(fa), "--batch", "-n", "50",
                              "-ms", "3", "--devices", dev], capture_output=True, text=True, timeout=300)
        assert out.returncode == 0, out.stderr
        assert out.stdout.splitlines() == lines, dev
