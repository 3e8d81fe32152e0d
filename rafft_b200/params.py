"""ViennaRNA energy-parameter files (`## RNAfold parameter file v2.0`) <-> struct rafft_energy_params.

The reference can load any ViennaRNA parameter file (rafft-rs/src/vienna.rs:75-81 `vrna_params_load`;
rufft `--params`); `read_par` fills the tables `rafft_create` takes, so a genuine `rna_turner2004.par`
replaces the built-in (reconstructed) tables entry by entry.  Only the 37 C free
energies are read (the `*_enthalpies` sections are skipped: this implementation folds at 37 C).

File layout as ViennaRNA's read_epars.c reads it: pair types CG GC GU UG AU UA NS (7), bases N A C G U (5);
every section is a flat list of integers in dcal/mol in C array order, `INF` = forbidden, `DEF` = default,
`/* ... */` comments anywhere.
  stack            7 x 7                      mismatch_*  7 x 5 x 5         dangle5/3  7 x 5
  int11            7 x 7 x 5 x 5              int21       7 x 7 x 5 x 5 x 5
  int22            6 x 6 x 4 x 4 x 4 x 4 (no NS pair, no N base)
  hairpin/bulge/interior  31 values           NINIO  m m_dH max             ML_params  cu cu_dH cc cc_dH ci ci_dH
  Misc  DuplexInit DuplexInit_dH TerminalAU TerminalAU_dH [LXC ...]         Triloops/Tetraloops/Hexaloops  SEQ E dH
`write_par` writes the same layout (NS / N rows filled with the neutral values ViennaRNA ships: 0 for
stacks and mismatches of NS, the maximum over the bases for N), so read_par(write_par(x)) == x."""
import ctypes as C
import re

from . import _native as N

INF = 10000000  # ViennaRNA energy_const.h
_MM = (("mismatch_hairpin", "mismatchH"), ("mismatch_interior", "mismatchI"), ("mismatch_interior_1n", "mismatch1nI"),
       ("mismatch_interior_23", "mismatch23I"), ("mismatch_multi", "mismatchM"), ("mismatch_exterior", "mismatchExt"))


def _tokens(text):
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    sections, cur = {}, None
    for line in text.splitlines():
        line = line.strip()
        if not line:
            continue
        if line.startswith("##"):
            continue
        if line.startswith("#"):
            name = line[1:].strip()
            if name.upper() == "END":
                break
            cur = sections.setdefault(name, [])
            continue
        if cur is not None:
            cur.extend(line.split())
    return sections


def _ints(tok, n, name, default=0):
    if len(tok) < n:
        raise ValueError("parameter file: section %s has %d values, %d expected" % (name, len(tok), n))
    out = []
    for t in tok[:n]:
        out.append(INF if t == "INF" else default if t == "DEF" else int(t))
    return out


def read_par(path):
    "ViennaRNA v2.0 parameter file -> _native.EnergyParams (37 C tables)"
    with open(path) as fh:
        text = fh.read()
    if "RNAfold parameter file v2.0" not in text.splitlines()[0]:
        raise ValueError("%s: not a '## RNAfold parameter file v2.0'" % path)
    S = _tokens(text)
    ep = N.EnergyParams()

    def need(name):
        if name not in S:
            raise ValueError("parameter file: section '# %s' is missing" % name)
        return S[name]
    v = _ints(need("stack"), 49, "stack")
    for a in range(6):
        for b in range(6):
            ep.stack[a * 6 + b] = v[a * 7 + b]
    for sec, field in _MM:
        v = _ints(need(sec), 7 * 25, sec)
        arr = getattr(ep, field)
        for t in range(6):
            for x in range(4):
                for y in range(4):
                    arr[(t * 4 + x) * 4 + y] = v[(t * 5 + x + 1) * 5 + y + 1]
    for sec, field in (("dangle5", "dangle5"), ("dangle3", "dangle3")):
        v = _ints(need(sec), 35, sec)
        arr = getattr(ep, field)
        for t in range(6):
            for x in range(4):
                arr[t * 4 + x] = v[t * 5 + x + 1]
    v = _ints(need("int11"), 49 * 25, "int11")
    for a in range(6):
        for b in range(6):
            for x in range(4):
                for y in range(4):
                    ep.int11[((a * 6 + b) * 4 + x) * 4 + y] = v[((a * 7 + b) * 5 + x + 1) * 5 + y + 1]
    v = _ints(need("int21"), 49 * 125, "int21")
    for a in range(6):
        for b in range(6):
            for x in range(4):
                for y in range(4):
                    for z in range(4):
                        ep.int21[(((a * 6 + b) * 4 + x) * 4 + y) * 4 + z] = \
                            v[(((a * 7 + b) * 5 + x + 1) * 5 + y + 1) * 5 + z + 1]
    v = _ints(need("int22"), 36 * 256, "int22")
    for k in range(36 * 256):
        ep.int22[k] = v[k]
    for sec in ("hairpin", "bulge", "interior"):
        v = _ints(need(sec), 31, sec)
        arr = getattr(ep, sec)
        for k in range(31):
            arr[k] = v[k]
    v = _ints(need("NINIO"), 3, "NINIO")
    ep.ninio, ep.max_ninio = v[0], v[2]
    v = _ints(need("ML_params"), 6, "ML_params")
    ep.ml_base, ep.ml_closing, ep.ml_intern = v[0], v[2], v[4]
    misc = need("Misc")
    ep.terminal_au = int(misc[2])
    ep.lxc = float(misc[4]) if len(misc) > 4 else 107.856
    for sec, field, width in (("Triloops", "tri", 5), ("Tetraloops", "tetra", 6), ("Hexaloops", "hexa", 8)):
        tok = S.get(sec, [])
        n = len(tok) // 3
        if n > N.MAX_SPECIAL:
            raise ValueError("parameter file: more than %d %s" % (N.MAX_SPECIAL, sec))
        setattr(ep, "n_" + field, n)
        for k in range(n):
            seq = tok[3 * k]
            if len(seq) != width:
                raise ValueError("parameter file: %s entry %r" % (sec, seq))
            getattr(ep, field)[k].value = seq.encode()
            getattr(ep, field + "_e")[k] = int(tok[3 * k + 1])
    return ep


def write_par(ep, path):
    "struct rafft_energy_params -> ViennaRNA v2.0 parameter file (free energies; enthalpy sections omitted)"
    out = ["## RNAfold parameter file v2.0", ""]

    def fmt(vals, per):
        return ["".join("%6s" % ("INF" if x >= INF else str(int(x))) for x in vals[k:k + per]) for k in range(0, len(vals), per)]
    st = [[ep.stack[a * 6 + b] if a < 6 and b < 6 else 0 for b in range(7)] for a in range(7)]
    out += ["# stack", "/*  CG     GC     GU     UG     AU     UA     @  */"] + fmt([x for r in st for x in r], 7) + [""]

    def with_n(get, nt, dims):
        "expand 6 pair types x 4 bases^dims to 7 x 5^dims: NS rows 0, N = max over the bases it stands for"
        import itertools
        vals = []
        for t in itertools.product(*[range(7)] * nt):
            for b in itertools.product(*[range(5)] * dims):
                if any(x == 6 for x in t):
                    vals.append(0)
                    continue
                choices = [range(4) if x == 0 else [x - 1] for x in b]
                vals.append(max(get(t, c) for c in itertools.product(*choices)))
        return vals
    for sec, field in _MM:
        arr = getattr(ep, field)
        out += ["# " + sec] + fmt(with_n(lambda t, c: arr[(t[0] * 4 + c[0]) * 4 + c[1]], 1, 2), 5) + [""]
    for sec in ("dangle5", "dangle3"):
        arr = getattr(ep, sec)
        out += ["# " + sec] + fmt(with_n(lambda t, c: arr[t[0] * 4 + c[0]], 1, 1), 5) + [""]
    out += ["# int11"] + fmt(with_n(lambda t, c: ep.int11[((t[0] * 6 + t[1]) * 4 + c[0]) * 4 + c[1]], 2, 2), 5) + [""]
    out += ["# int21"] + fmt(with_n(lambda t, c: ep.int21[(((t[0] * 6 + t[1]) * 4 + c[0]) * 4 + c[1]) * 4 + c[2]], 2, 3), 5) + [""]
    out += ["# int22"] + fmt(list(ep.int22), 4) + [""]
    for sec in ("hairpin", "bulge", "interior"):
        out += ["# " + sec] + fmt(list(getattr(ep, sec)), 10) + [""]
    out += ["# NINIO", "/* Ninio = MIN(max, m*|n1-n2| */", "/*\t    m\t  m_dH     max  */", "\t%d\t320\t%d" % (ep.ninio, ep.max_ninio), ""]
    out += ["# ML_params", "/* F = cu*n_unpaired + cc + ci*loop_degree (+TermAU) */", "/*\t    cu\t    cu_dH\t    cc\t    cc_dH\t    ci\t    ci_dH  */",
            "\t%d\t0\t%d\t3000\t%d\t-220" % (ep.ml_base, ep.ml_closing, ep.ml_intern), ""]
    out += ["# Misc", "/* all parameters are pseudo energies in the SAME units as the others */",
            "/*      DuplexInit     TerminalAU   LXC  */", "\t410\t360\t%d\t370\t%f\t0" % (ep.terminal_au, ep.lxc), ""]
    for sec, field in (("Triloops", "tri"), ("Tetraloops", "tetra"), ("Hexaloops", "hexa")):
        out += ["# " + sec]
        for k in range(getattr(ep, "n_" + field)):
            out.append("%s %6d %6d" % (getattr(ep, field)[k].value.decode(), getattr(ep, field + "_e")[k], 0))
        out += [""]
    out += ["#END", ""]
    with open(path, "w") as fh:
        fh.write("\n".join(out))
