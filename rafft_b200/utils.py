"""Host-side types and text helpers of the folding path, mirroring the names of the
reference's rafft/utils.py (Structure :32-39, dot_bracket :42-50, paired_positions :53-67,
read_fasta :161-169, parse_rafft_output :172-185).  The numeric kernels of that file
(prep_sequence, auto_cor, window_slide, eval_one_struct, get_inner_loop/get_outer_loop) have
no host counterpart here: they run inside the CUDA library."""
import struct as _struct


def energy_from_dcal(e_dcal):
    """ViennaRNA returns `(float)energy / 100.` (float32) which Python widens to double
    (SURVEY.md Q9); printing with %6.1f must start from that value."""
    return _struct.unpack("f", _struct.pack("f", e_dcal / 100.0))[0]


class Structure:
    """A saved structure.  Consumers of the reference read `.str_struct` and `.energy`
    (bin/rafft:62-79, rafft/rafft_kin.py:50,76,110-115); `.pair_list` is derived on demand.
    `.energy_dcal` is the exact integer energy (dcal/mol)."""

    __slots__ = ("node_list", "energy", "energy_dcal", "str_struct", "_pairs")

    def __init__(self, node_list=None, pair_list=None):
        self.node_list = node_list if node_list is not None else []
        self.energy = 0.0
        self.energy_dcal = 0
        self.str_struct = ""
        self._pairs = pair_list

    @property
    def pair_list(self):
        if self._pairs is None:
            self._pairs = paired_positions(self.str_struct)
        return self._pairs

    @pair_list.setter
    def pair_list(self, value):
        self._pairs = value

    def __repr__(self):
        return "Structure(%r, %.2f)" % (self.str_struct, self.energy)


def dot_bracket(pair_list, len_seq, SEQ=None):
    "list of base pairs -> dot-bracket text"
    out = ["."] * len_seq
    for i, j in pair_list:
        out[i] = "("
        out[j] = ")"
    return "".join(out)


def paired_positions(structure):
    "dot-bracket text (also <> and [] pseudo-knot brackets) -> list of (i, j)"
    openers = {"(": 0, "<": 0, "[": 1}
    closers = {")": 0, ">": 0, "]": 1}
    stacks = ([], [])
    pairs = []
    for k, ch in enumerate(structure):
        if ch in openers:
            stacks[openers[ch]].append(k)
        elif ch in closers:
            pairs.append((stacks[closers[ch]].pop(), k))
    return pairs


def read_fasta(infile):
    "FASTA file -> {name: sequence}"
    out = {}
    name = None
    with open(infile) as fh:
        for line in fh:
            if line.startswith(">"):
                name = line.strip()[1:]
                out[name] = ""
            else:
                out[name] += line.strip()
    return out


def parse_rafft_output(infile):
    """`rafft --traj` text -> (list of steps, each a list of Structure; sequence)."""
    steps = []
    with open(infile) as fh:
        seq = fh.readline().strip()
        for line in fh:
            if line.startswith("# --"):
                steps.append([])
            else:
                db, nrj = line.strip().split()
                st = Structure()
                st.str_struct = db
                st.energy = float(nrj)
                st.energy_dcal = int(round(st.energy * 100))
                steps[-1].append(st)
    return steps, seq


def format_trajectory(sequence, trajectory):
    "the text `rafft --traj` prints (bin/rafft:71-79)"
    lines = [sequence]
    for si, step in enumerate(trajectory):
        lines.append("# {:-^20}".format(si))
        for st in step:
            lines.append("%s %6.1f" % (st.str_struct, st.energy))
    return "\n".join(lines) + "\n"
