"""Host-side logic: reference-compatible types/text formats, CLI argument surface, synthetic
inputs, sharding (world_size 2 over gloo).  CPU only."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
from conftest import GOLDEN  # noqa: E402


def test_energy_print_rounding():
    "float32(E/100) then %6.1f (SURVEY Q9): -0.35 prints -0.3, -0.15 prints -0.2"
    from rafft_b200.utils import energy_from_dcal
    assert "%6.1f" % energy_from_dcal(-35) == "  -0.3"
    assert "%6.1f" % energy_from_dcal(-15) == "  -0.2"
    assert "%6.1f" % energy_from_dcal(-1090) == " -10.9"


def test_parse_and_format_roundtrip():
    from rafft_b200.utils import format_trajectory, parse_rafft_output
    for name in ("rafft.out", "rafft_20.out"):
        path = os.path.join(GOLDEN, name)
        steps, seq = parse_rafft_output(path)
        assert format_trajectory(seq, steps) == open(path).read()
        assert all(len(s.str_struct) == len(seq) for st in steps for s in st)


def test_pairs_and_dot_bracket():
    from rafft_b200.utils import Structure, dot_bracket, paired_positions
    db = "((..((...))..))."
    pl = paired_positions(db)
    assert sorted(pl) == [(0, 14), (1, 13), (4, 10), (5, 9)]
    assert dot_bracket(pl, len(db)) == db
    s = Structure()
    s.str_struct = db
    assert sorted(s.pair_list) == sorted(pl)
    assert paired_positions("<[>]") == [(0, 2), (1, 3)]


def test_fold_signature_matches_reference():
    import inspect
    from rafft_b200 import fold
    sig = inspect.signature(fold)
    assert list(sig.parameters) == ["sequence", "nb_mode", "max_stack", "max_branch", "min_hp", "min_nrj",
                                    "traj", "temp", "gc_wei", "au_wei", "gu_wei"]
    d = {k: v.default for k, v in sig.parameters.items() if v.default is not inspect.Parameter.empty}
    assert d == dict(nb_mode=100, max_stack=1, max_branch=100, min_hp=3, min_nrj=0.0, traj=False, temp=37.0,
                     gc_wei=3.0, au_wei=2.0, gu_wei=1.0)


def test_invalid_letter_is_keyerror():
    from rafft_b200 import fold
    with pytest.raises(KeyError):
        fold("GGGAXACCC")


def test_cli_flags():
    import importlib.machinery
    import importlib.util
    loader = importlib.machinery.SourceFileLoader("rafft_cli", os.path.join(ROOT, "bin", "rafft"))
    spec = importlib.util.spec_from_loader("rafft_cli", loader)
    cli = importlib.util.module_from_spec(spec)
    loader.exec_module(cli)
    a = cli.parse_arguments(["-s", "GGGAAACCC"])
    assert (a.n_mode, a.max_stack, a.max_branch, a.min_hp, a.min_nrj, a.temp) == (100, 1, 1000, 3, 0, 37.0)
    assert (a.gc_wei, a.au_wei, a.gu_wei, a.traj, a.bench) == (3.0, 2.0, 1.0, False, False)
    a = cli.parse_arguments(["-s", "G", "-n", "200", "-ms", "200", "-tr", "--bench", "-mb", "2", "-p", "2", "--bp_only"])
    assert (a.n_mode, a.max_stack, a.traj, a.bench) == (200, 200, True, True)


def test_synthetic_inputs():
    from rafft_b200.synth import benchmark_lengths, config_batch, splitmix_seq
    assert splitmix_seq(1, 120).startswith("GGUCCUUGCUCGCGCAGUGU")
    pool = benchmark_lengths()
    assert len(pool) == 2248 and pool[0] >= 50 and pool[-1] <= 500 and pool == sorted(pool)
    b = config_batch("cfg2", 300)
    assert b == config_batch("cfg2", 300) and b[:100] == config_batch("cfg2", 100)
    assert all(50 <= len(s) <= 500 and set(s) <= set("ACGU") for s in b)
    assert [len(s) for s in config_batch("cfg4", 2)] == [3000, 3000]


def test_partition_properties():
    from rafft_b200.shard import imbalance, partition
    from rafft_b200.synth import config_batch
    lens = [len(s) for s in config_batch("cfg2", 2000)]
    for w in (1, 2, 4, 8):
        sh = partition(lens, w, 50)
        allidx = np.concatenate(sh)
        assert sorted(allidx.tolist()) == list(range(len(lens)))
        assert imbalance(lens, sh, 50) < 1.02
    assert [s.tolist() for s in partition([], 2)] == [[], []]
    assert sum(len(s) for s in partition([10], 4)) == 1


_WORKER = r'''
import os, sys
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "oracle"))
import torch.distributed as dist
import pyoracle
from rafft_b200.shard import fold_batch_sharded
from rafft_b200.synth import config_batch
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
seqs = config_batch("cfg2", 24)
calls = []
def fold_fn(shard):                      # the oracle stands in for the GPU fold on CPU
    calls.append(len(shard))
    return pyoracle.fold_batch(shard, 20, 5, 100, nthreads=2)[0]
out = fold_batch_sharded(seqs, fold_fn, rank, world, max_stack=5)
if rank == 0:
    want = pyoracle.fold_batch(seqs, 20, 5, 100, nthreads=2)[0]
    assert out == want
    print("SHARD_OK", world, calls[0])
else:
    assert out is None
dist.barrier(); dist.destroy_process_group()
'''


def test_sharded_fold_world2_gloo(tmp_path, oracle):
    script = tmp_path / "w.py"
    script.write_text(_WORKER % {"root": ROOT})
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", str(script)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "SHARD_OK 2" in r.stdout
