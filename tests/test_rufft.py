"""The rafft-rs surface (SURVEY.md 8(f) "next" #2; reference rafft-rs/src/bindings.rs:11-81,
rafft-rs/src/bin/rufft.rs:96-183) over the Python-semantics engine.

not gpu: graph assembly (numbering, depths, edges, rates) from oracle trajectories + the oracle's
         connectivity; the Rust twin's own known-answer test (fast_folding.rs:108-126) on the oracle;
         text formats; CLI flags.
gpu:     FastFoldingGraph / bin/rufft through the C-ABI against the same assembly fed by the oracle."""
import math
import os
import subprocess
import sys

import pytest

from conftest import ROOT
from rafft_b200.synth import splitmix_seq

RUST_KAT_SEQ = "GGGUUUGCGGUGUAAGUGCAGCCCGUCUUACACCGUGCGGCACAGGCACUAGUACUGAUGUCGUAUACAGGGCUUUUGACAU"
RUST_KAT_END = "..((((((((((((((.((.....))))))))))))).))).(((.........)))((((((.............))))))"


def oracle_graph(oracle, seq, lags=100, branches=1000, saved=1):
    import kin_oracle
    from rafft_b200.rufft import graph_from_trajectory
    steps = oracle.fold(seq, lags, saved, branches)
    uniq, conn = kin_oracle.folding_graph([[db for db, _ in st] for st in steps])
    return graph_from_trajectory(steps, uniq, conn), steps


def test_rust_known_answer_on_oracle(oracle):
    "fast_folding.rs:108-126: default config, one saved trajectory -> the last structure inserted"
    (nodes, edges, steps), _ = oracle_graph(oracle, RUST_KAT_SEQ)
    assert nodes[-1][1] == RUST_KAT_END
    assert nodes[0] == (0, "." * len(RUST_KAT_SEQ), 0)
    assert [d for d, _, _ in nodes] == list(range(len(nodes)))      # one new structure per level
    assert edges == [(i, i + 1) for i in range(len(nodes) - 1)]     # a path
    assert steps == [[k] for k in range(len(nodes))]


@pytest.mark.parametrize("saved", [1, 5, 20])
def test_graph_assembly_properties(oracle, saved):
    import kin_oracle
    seq = splitmix_seq(77, 90)
    (nodes, edges, steps), traj = oracle_graph(oracle, seq, 50, 100, saved)
    dbs = [db for _, db, _ in nodes]
    assert len(set(dbs)) == len(dbs)
    # insertion order = first appearance, depth = step of first appearance
    seen = []
    for t, st in enumerate(traj):
        for db, e in st:
            if db not in seen:
                seen.append(db)
                assert nodes[len(seen) - 1] == (t, db, e)
    assert seen == dbs
    assert [[dbs[k] for k in st] for st in steps] == [[db for db, _ in st] for st in traj]
    assert len(set(edges)) == len(edges)
    pairs = [set(kin_oracle.paired_positions(db)) for db in dbs]
    for i, j in edges:
        assert i != j and pairs[i] < pairs[j]                     # strictly nested pair sets
        assert nodes[j][0] >= 1
    for j in range(1, len(nodes)):                                 # everything but the root has a parent
        assert any(b == j for _, b in edges)


def test_transition_rates_layout(oracle):
    "folding_graph.rs:143-178: diagonal entries first, then rate / reverse rate per edge, in dcal"
    from rafft_b200.rufft import FastFoldingGraph
    (nodes, edges, steps), _ = oracle_graph(oracle, RUST_KAT_SEQ, saved=5)
    g = FastFoldingGraph(RUST_KAT_SEQ, saved_trajectories=5)
    assert g.trajectories.__doc__ and g.directed_edges() == ([], [])     # root only before construction
    assert g.transition_rates() == ([0.0], [0], [0])
    g._nodes, g._edges, g._steps, g._built = nodes, edges, steps, True    # host logic only: no device here
    data, iis, jjs = g.transition_rates(0.61)
    n = len(nodes)
    assert len(data) == n + 2 * len(edges) and iis[:n] == jjs[:n] == list(range(n))
    row = [0.0] * n
    for k, (i, j) in enumerate(edges):
        ddg = float(nodes[j][2] - nodes[i][2])
        r = min(1.0, math.exp(-0.61 * ddg)) if -0.61 * ddg < 700 else 1.0
        rr = min(1.0, math.exp(0.61 * ddg)) if 0.61 * ddg < 700 else 1.0
        assert (data[n + 2 * k], iis[n + 2 * k], jjs[n + 2 * k]) == (r, i, j)
        assert (data[n + 2 * k + 1], iis[n + 2 * k + 1], jjs[n + 2 * k + 1]) == (rr, j, i)
        row[i] -= r
        row[j] -= rr
    assert data[:n] == row
    assert g.directed_edges() == ([i for i, _ in edges], [j for _, j in edges])


def test_text_formats(oracle):
    from rafft_b200 import rufft as R
    (nodes, edges, steps), _ = oracle_graph(oracle, RUST_KAT_SEQ)
    g = R.FastFoldingGraph(RUST_KAT_SEQ)
    g._nodes, g._edges, g._steps, g._built = nodes, edges, steps, True
    lines = R.format_default(g).splitlines()
    assert lines[0] == "[0] %s 0.00" % ("." * len(RUST_KAT_SEQ))
    assert lines[-1] == "[%d] %s %.2f" % (len(nodes) - 1, RUST_KAT_END, nodes[-1][2] * 0.01)
    compat = R.format_compat(g).splitlines()
    assert compat[0] == "# ---------0----------" and compat[2] == "# ---------1----------"
    assert compat[-1].split()[0] == RUST_KAT_END
    b = R.format_benchmark(g).split()
    assert b[0] == RUST_KAT_SEQ and b[1] == "82" and b[2] == RUST_KAT_END
    assert b[3] == "%.1f" % (nodes[-1][2] * 0.01) and int(b[4]) == RUST_KAT_END.count("(")
    assert R.format_edges(g) == "".join("%d %d\n" % e for e in edges)


def test_invalid_sequence_is_rejected():
    from rafft_b200.rufft import FastFoldingGraph
    for bad in ("GGGNAAACCC", "gggaaaccc", "GGGTAAACCC"):     # encoding.rs:104-143 accepts A C G U only
        with pytest.raises(ValueError, match="Not a valid RNA Sequence"):
            FastFoldingGraph(bad)


def test_rufft_cli_flags():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bin", "rufft"), "--help"], capture_output=True, text=True)
    assert out.returncode == 0
    for flag in ("--params", "--AU", "--GC", "--GU", "--temperature", "--min-unpaired", "--minimum-helix-energy",
                 "--positional-lags", "--branch", "--saved-trajectories", "--benchmark", "--compat", "--output-edges"):
        assert flag in out.stdout


# ------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_fast_folding_graph_rust_kat(ctx):
    from rafft_b200.rufft import FastFoldingGraph
    g = FastFoldingGraph(RUST_KAT_SEQ)                        # defaults of bindings.rs:19-28
    traj = g.trajectories()
    assert traj[-1][1] == RUST_KAT_END
    assert traj[-1][0] == len(traj) - 1 and abs(traj[-1][2] - (-20.5)) < 1e-12
    assert g.directed_edges() == (list(range(len(traj) - 1)), list(range(1, len(traj))))


@pytest.mark.gpu
@pytest.mark.parametrize("seed,length,lags,saved", [(5, 60, 100, 1), (77, 90, 50, 5), (9, 150, 100, 20),
                                                    (200, 200, 100, 50)])
def test_fast_folding_graph_vs_oracle(ctx, oracle, seed, length, lags, saved):
    from rafft_b200.rufft import FastFoldingGraph
    seq = splitmix_seq(seed, length)
    (nodes, edges, steps), _ = oracle_graph(oracle, seq, lags, 1000, saved)
    g = FastFoldingGraph(seq, number_of_lags=lags, saved_trajectories=saved)
    assert g.trajectories() == [(d, db, e * 0.01) for d, db, e in nodes]
    assert g.directed_edges() == ([i for i, _ in edges], [j for _, j in edges])
    assert g.steps() == steps
    data, iis, jjs = g.transition_rates()
    assert len(data) == len(nodes) + 2 * len(edges)
    rows = [0.0] * len(nodes)
    for v, i, j in zip(data, iis, jjs):
        rows[i] += v
    assert all(abs(r) < 1e-9 for r in rows)                  # generator matrix: rows sum to zero


@pytest.mark.gpu
def test_rufft_cli(ctx, oracle, tmp_path):
    from rafft_b200 import rufft as R
    seq = splitmix_seq(77, 90)
    (nodes, edges, steps), _ = oracle_graph(oracle, seq, 50, 1000, 5)
    exe = [sys.executable, os.path.join(ROOT, "bin", "rufft"), seq, "-l", "50", "-s", "5"]
    efile = str(tmp_path / "edges.txt")
    out = subprocess.run(exe + ["-o", efile], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert out.stdout == "".join("[%d] %s %.2f\n" % (d, db, e * 0.01) for d, db, e in nodes)
    assert open(efile).read() == "".join("%d %d\n" % e for e in edges)
    out = subprocess.run(exe + ["--compat"], capture_output=True, text=True)
    want = "".join("# ---------%d----------\n" % t + "".join("%s %.2f\n" % (nodes[k][1], nodes[k][2] * 0.01) for k in st)
                   for t, st in enumerate(steps))
    assert out.stdout == want
    # the blocks parse with the reference-format reader of the kinetics (rafft/utils.py:172-185)
    from rafft_b200.utils import parse_rafft_output
    p = tmp_path / "compat.out"
    p.write_text(seq + "\n" + out.stdout)
    parsed = parse_rafft_output(str(p))
    assert [len(s) for s in parsed[0]] == [len(s) for s in steps]
    out = subprocess.run(exe + ["-B"], capture_output=True, text=True)
    best = sorted(nodes, key=lambda nd: nd[2])[:5]
    assert out.stdout == "".join("%s %d %s %.1f %d\n" % (seq, len(seq), db, e * 0.01, db.count("(")) for _, db, e in best)
    assert subprocess.run(exe + ["-T", "25"], capture_output=True, text=True).returncode != 0
    assert R.FastFoldingGraph(seq).trajectories()[0][1] == "." * len(seq)
