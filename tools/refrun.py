"""Run the UNMODIFIED reference Python (/root/reference) through oracle/shim.  Tooling for
golden-vector generation and oracle validation; only usable in the build container."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, "/root/reference")
sys.path.insert(0, os.path.join(ROOT, "oracle", "shim"))


def splitmix_seq(seed, length):
    """SURVEY.md section 8(d) generator: base = "ACGU"[splitmix64 >> 62]."""
    M = (1 << 64) - 1
    s = seed & M
    out = []
    for _ in range(length):
        s = (s + 0x9E3779B97F4A7C15) & M
        z = s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M
        z ^= z >> 31
        out.append("ACGU"[z >> 62])
    return "".join(out)


def ref_fold(seq, nb_mode=100, max_stack=1, max_branch=1000, min_hp=3, min_nrj=0.0,
             gc=3.0, au=2.0, gu=1.0):
    """-> list of steps, each a list of (dot-bracket, E_dcal)."""
    from rafft.rafft import fold
    _, traj = fold(seq, nb_mode, max_stack, max_branch, min_hp, min_nrj, True, 37.0, gc, au, gu)
    return [[(s.str_struct, int(round(s.energy * 100))) for s in step] for step in traj]
