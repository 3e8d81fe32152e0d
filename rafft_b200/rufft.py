"""The rafft-rs surface over the GPU engine: `FastFoldingGraph` with the
constructor arguments, defaults and methods of the pyo3 class (rafft-rs/src/bindings.rs:11-81)
and the text formats of `rufft` (rafft-rs/src/bin/rufft.rs:120-183).

The engine underneath is the one of `rafft_b200.rafft.fold`, i.e. the PYTHON reference's
beam search (rafft/rafft.py:156-216), which is the normative semantics of this repository.  The
Rust twin's breadth-first search differs from it in documented ways (per-step `seen`, only new
structures expand, lag ties in ascending index, f64 FFT noise in the ranking: SURVEY Note R), so
for `saved_trajectories` > 1 the structure lists of the two can differ; for
`saved_trajectories` = 1 and an exact correlation they coincide, and the Rust twin's own
known-answer test (rafft-rs/src/fast_folding.rs:108-126) is reproduced (tests/test_rufft.py).

How the graph is read off the trajectory (all of it on the device except the numbering):
  nodes   unique structures in order of first appearance (step by step, beam order = energy
          order), the iteration order of RafftGraph::iter (folding_graph.rs:131-133);
  depth   the folding step at which the structure first entered the beam (folding_graph.rs:93);
  edges   (i -> j) for structures of consecutive steps whose pair sets are nested (every pair of
          i is a pair of j, i != j): the connectivity of the reference's own kinetics
          (rafft/rafft_kin.py:48-56), computed by `rafft_kin_graph`; each edge once, as
          DiGraph::update_edge keeps them (folding_graph.rs:116).
          DIVERGENCE from rafft-rs for `saved_trajectories` > 1: the Rust twin adds one edge
          parent -> child per child it inserts (folding_graph.rs insert -> update_edge(parent,
          node)); the nested-pair-set rule used here (the Python reference's kinetics rule) also
          links an older beam member A, retained beside its child B, to B's child C (A's pairs are
          a subset of C's), so `transition_rates()`, `directed_edges()` and `rufft -o` can carry
          extra transitive edges there.  With `saved_trajectories` = 1 (the default and the only
          setting the twin's own test pins) a step has one structure and the two rules coincide.
"""
import math

import numpy as np

from .rafft import fold_batch, get_context
from .rafft_kin import folding_graph

_VALID = frozenset("ACGU")


def graph_from_trajectory(steps, uniq, conn):
    """steps: folding steps of (dot-bracket, E_dcal); uniq[r] = first row equal to row r of the
    concatenated steps; conn[t][c] = indices (in step t-1) of the structures nested in structure
    c of step t -- the outputs of `rafft_kin_graph`.
    -> (nodes [(depth, dot-bracket, E_dcal)], edges [(i, j)], per-step node indices)."""
    flat = [(t, db, e) for t, step in enumerate(steps) for db, e in step]
    index, nodes = {}, []
    for r, (t, db, e) in enumerate(flat):
        if uniq[r] == r:
            index[r] = len(nodes)
            nodes.append((t, db, int(e)))
    node_of = [index[uniq[r]] for r in range(len(flat))]
    off = [0]
    for step in steps:
        off.append(off[-1] + len(step))
    edges, have = [], set()
    for t in range(1, len(steps)):  # step 0 is the root: nothing precedes it
        for c, prevs in enumerate(conn[t]):
            j = node_of[off[t] + c]
            for p in prevs:
                i = node_of[off[t - 1] + p]
                if i != j and (i, j) not in have:
                    have.add((i, j))
                    edges.append((i, j))
    return nodes, edges, [[node_of[r] for r in range(off[t], off[t + 1])] for t in range(len(steps))]


class FastFoldingGraph:
    """Drop-in for `rafft.FastFoldingGraph` of the Rust twin (bindings.rs:11-81)."""

    def __init__(self, sequence, number_of_lags=100, number_of_branches=1000, saved_trajectories=1,
                 au=2.0, gc=3.0, gu=1.0, min_unpaired=3, min_loop_energy=0.0, device=0, parameters=None):
        for ch in sequence:
            if ch not in _VALID:
                # EncodedSequence::with_basepair_weights -> Err(InvalidNucleotide), which
                # folding_graph() turns into this panic message (fast_folding.rs:91-92)
                raise ValueError("Not a valid RNA Sequence!: InvalidNucleotide(%r)" % ch)
        if len(sequence) == 0:
            raise ValueError("empty sequence")
        self.sequence = sequence
        self.number_of_lags = int(number_of_lags)
        self.number_of_branches = int(number_of_branches)
        self.saved_trajectories = int(saved_trajectories)
        self.au, self.gc, self.gu = float(au), float(gc), float(gu)
        self.min_unpaired = int(min_unpaired)
        self.min_loop_energy = float(min_loop_energy)
        self.device = int(device)
        # a ViennaRNA parameter file (rufft --params, vienna.rs:75-81 vrna_params_load) replaces the built-in tables
        self.parameters = parameters
        # RafftGraph::new holds only the root (folding_graph.rs:38-71)
        self._nodes = [(0, "." * len(sequence), 0)]  # (depth, dot-bracket, E in dcal/mol)
        self._edges = []
        self._steps = [[0]]
        self._built = False

    # -- construction --------------------------------------------------------------------
    def construct_trajectories(self):
        """RafftGraph::construct_trajectories (folding_graph.rs:124-127): one GPU fold with the
        full trajectory, then the folding graph of that trajectory."""
        if self._built:
            return
        if self.parameters is not None:
            from .params import read_par
            from .rafft import FoldContext
            ctx = FoldContext(self.device, read_par(self.parameters))
        else:
            ctx = get_context(self.device)
        _, trajectory = fold_batch([self.sequence], self.number_of_lags, self.saved_trajectories,
                                   self.number_of_branches, self.min_unpaired, self.min_loop_energy,
                                   True, 37.0, self.gc, self.au, self.gu, context=ctx)[0]
        steps = [[(st.str_struct, st.energy_dcal) for st in step] for step in trajectory]
        uniq, conn, conn_off, step_off = folding_graph(trajectory, context=ctx)
        blocks = []
        for t in range(len(steps)):
            n_cur, n_prev = len(steps[t]), len(steps[t - 1])
            blk = conn[int(conn_off[t]):int(conn_off[t + 1])].reshape(n_cur, n_prev)
            blocks.append([np.nonzero(row)[0].tolist() for row in blk])
        self._nodes, self._edges, self._steps = graph_from_trajectory(steps, list(map(int, uniq)), blocks)
        self._built = True

    # -- the pyo3 methods ----------------------------------------------------------------
    def trajectories(self):
        "-> [(depth, dot-bracket, kcal/mol)] in insertion order (bindings.rs:54-70)"
        self.construct_trajectories()
        return [(d, db, e * 0.01) for d, db, e in self._nodes]

    def transition_rates(self, beta=0.61):
        """COO triplets (values, is, js) of the Metropolis rates, laid out as
        RafftGraph::transition_rates does (folding_graph.rs:143-178): the diagonal first, then
        r(i->j), r(j->i) per edge.  As there, the energy difference enters in dcal/mol."""
        n = len(self._nodes)
        data = [0.0] * n
        iis = list(range(n))
        jjs = list(range(n))
        for i, j in self._edges:
            ddg = float(self._nodes[j][2] - self._nodes[i][2])
            rate = min(1.0, _exp(-beta * ddg))
            rev_rate = min(1.0, _exp(beta * ddg))
            data.append(rate); iis.append(i); jjs.append(j)
            data.append(rev_rate); iis.append(j); jjs.append(i)
            data[i] -= rate
            data[j] -= rev_rate
        return data, iis, jjs

    def directed_edges(self):
        "-> (is, js) of the directed edges (bindings.rs:77-80)"
        return [i for i, _ in self._edges], [j for _, j in self._edges]

    # -- what rufft prints ---------------------------------------------------------------
    def steps(self):
        "node indices of the beam of every folding step (with the duplicates `--compat` prints)"
        self.construct_trajectories()
        return [list(s) for s in self._steps]


def _exp(x):
    try:
        return math.exp(x)
    except OverflowError:  # f64::exp saturates to +inf
        return math.inf


def format_default(graph):
    "`[depth] structure energy` per structure (rufft.rs:121-130)"
    return "".join("[%d] %s %.2f\n" % (d, db, e) for d, db, e in graph.trajectories())


def format_compat(graph):
    """`--compat`: the blocks the kinetics script parses, duplicates included (rufft.rs:131-160).
    The Rust twin has to rebuild the beams from its node list; here they are the beams."""
    nodes = graph.trajectories()
    out = []
    for depth, step in enumerate(graph.steps()):
        out.append("# ---------%d----------\n" % depth)
        for k in step:
            out.append("%s %.2f\n" % (nodes[k][1], nodes[k][2]))
    return "".join(out)


def format_benchmark(graph):
    "`--benchmark`: `seq len structure energy pairs` of the best structures (rufft.rs:169-182)"
    nodes = sorted(graph.trajectories(), key=lambda nd: nd[2])  # stable, as sort_by_key
    seq = graph.sequence
    return "".join("%s %d %s %.1f %d\n" % (seq, len(seq), db, e, db.count("("))
                   for _, db, e in nodes[:min(graph.saved_trajectories, len(nodes))])


def format_edges(graph):
    "`-o FILE`: one `i j` line per directed edge (rufft.rs:162-168)"
    return "".join("%d %d\n" % (i, j) for i, j in zip(*graph.directed_edges()))


__all__ = ["FastFoldingGraph", "format_default", "format_compat", "format_benchmark", "format_edges"]
