"""Landscape inputs of the reference's utility/surface.py: the all-pairs base-pair distance matrix
of the unique structures of a folding trajectory (get_distance_matrix, utility/surface.py:19-26),
computed on the GPU through `rafft_bp_distance_matrix` (include/rafft_b200.h), and the parser that
feeds it (parse_rafft_output, :29-40: unique structures in order of first appearance).

The reference then embeds the matrix with sklearn MDS (:97-100) and draws it with matplotlib;
`mds_coordinates` mirrors the embedding call, the drawing is unavailable on this box."""
import ctypes as C

import numpy as np

from . import _native as N
from .rafft import get_context


def parse_rafft_output(infile):
    """`rafft --traj` text -> ([(structure, energy)] without repeats, sequence)."""
    results, seen = [], set()
    with open(infile) as fh:
        seq = fh.readline().strip()
        for line in fh:
            if not line.startswith("#"):
                struct, nrj = line.strip().split()
                if struct not in seen:
                    results.append((struct, float(nrj)))
                    seen.add(struct)
    return results, seq


def get_distance_matrix(structures, device=0, context=None):
    """structures: [(dot-bracket, energy)] -> float64 matrix of base-pair distances (the
    reference fills a numpy.zeros matrix with RNA.bp_distance, utility/surface.py:20-25)."""
    ctx = context if context is not None else get_context(device)
    rows = [s for s, _ in structures]
    n = len(rows)
    if n == 0:
        return np.zeros((0, 0))
    L = len(rows[0])
    if any(len(r) != L for r in rows):
        raise ValueError("structures of one landscape must have the same length")
    dist = np.zeros((n, n), dtype=np.int32)
    rc = ctx._lib.rafft_bp_distance_matrix(ctx._h, "".join(rows).encode(), n, L,
                                           dist.ctypes.data_as(C.POINTER(C.c_int32)))
    if rc == N.RAFFT_ERR_SEQ:
        raise ValueError("unbalanced brackets")
    if rc != N.RAFFT_OK:
        raise ctx._err(rc)
    return dist.astype(np.float64)


def mds_coordinates(dist_mat, seed=3):
    "the 2-D embedding surface.py draws (utility/surface.py:96-100); needs scikit-learn"
    from numpy.random import RandomState
    from sklearn import manifold
    mds = manifold.MDS(n_components=2, max_iter=5000, eps=1e-9, random_state=RandomState(seed=seed),
                       dissimilarity="precomputed")
    return mds.fit_transform(dist_mat)
