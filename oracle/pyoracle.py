"""ORACLE access from Python -- test infrastructure, NOT the product.

ctypes wrapper of oracle/_ref/liboracle.so (turner_eval.c + rafft_oracle.cpp).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this."""
import ctypes as C
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB_PATH = os.path.join(HERE, "_ref", "liboracle.so")
sys.path.insert(0, ROOT)
from rafft_b200._native import CNT_N, EnergyParams, Params  # noqa: E402  (struct layouts only)


def build(force=False):
    srcs = [os.path.join(HERE, f) for f in ("turner_eval.c", "rafft_oracle.cpp", "Makefile")]
    srcs += [os.path.join(ROOT, "include", "rafft_b200.h"), os.path.join(ROOT, "rafft_b200", "csrc", "turner2004.inc")]
    if not force and os.path.exists(LIB_PATH) and all(os.path.getmtime(s) <= os.path.getmtime(LIB_PATH) for s in srcs):
        return LIB_PATH
    r = subprocess.run(["make", "-C", HERE, "-B"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        L.oracle_default_params.argtypes = [C.POINTER(EnergyParams)]
        L.oracle_eval_db.argtypes = [C.POINTER(EnergyParams), C.c_char_p, C.c_char_p, C.c_int]
        L.oracle_eval_db.restype = C.c_int
        L.oracle_eval_pt.argtypes = [C.POINTER(EnergyParams), C.c_char_p, C.POINTER(C.c_int), C.c_int]
        L.oracle_eval_pt.restype = C.c_int
        L.oracle_fold.argtypes = [C.POINTER(EnergyParams), C.c_char_p, C.c_int, C.POINTER(Params),
                                  C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
        L.oracle_fold.restype = C.c_int
        L.oracle_free.argtypes = [C.c_void_p]
        L.oracle_fold_batch.argtypes = [C.POINTER(EnergyParams), C.c_char_p, C.POINTER(C.c_int64), C.c_int,
                                        C.POINTER(Params), C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int),
                                        C.POINTER(C.c_int64)]
        L.oracle_fold_batch.restype = C.c_int
        L.oracle_fold_batch_check.argtypes = [C.POINTER(EnergyParams), C.c_char_p, C.POINTER(C.c_int64), C.c_int,
                                              C.POINTER(Params), C.c_int, C.c_int] + [C.c_void_p] * 10
        L.oracle_fold_batch_check.restype = C.c_int
        L.oracle_auto_cor.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_double, C.c_double,
                                      C.c_double, C.POINTER(C.c_double)]
        L.oracle_window_slide.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_int, C.c_int,
                                          C.c_double, C.c_double, C.c_double, C.POINTER(C.c_int),
                                          C.POINTER(C.c_double)]
        _lib = L
    return _lib


_P = None


def default_params():
    global _P
    if _P is None:
        _P = EnergyParams()
        lib().oracle_default_params(C.byref(_P))
    return _P


def eval_db(seq, db, params=None):
    return lib().oracle_eval_db(C.byref(params or default_params()), seq.encode(), db.encode(), len(seq))


def parse_listing(text):
    """'# k' / '<db> <E>' listing -> list of steps of (db, E_dcal)."""
    steps = []
    for line in text.splitlines():
        if line.startswith("#"):
            steps.append([])
        elif line:
            db, e = line.split()
            steps[-1].append((db, int(e)))
    return steps


def fold(seq, nb_mode=100, max_stack=1, max_branch=1000, min_hp=3, min_nrj=0.0, gc=3.0, au=2.0, gu=1.0,
         counters=False):
    """-> list of steps, each a list of (dot-bracket, E_dcal) (the --traj order)."""
    prm = Params(nb_mode, max_stack, max_branch, min_hp, 1, min_nrj, 37.0, gc, au, gu)
    txt = C.c_void_p()
    cnt = (C.c_int64 * CNT_N)()
    rc = lib().oracle_fold(C.byref(default_params()), seq.encode(), len(seq), C.byref(prm), C.byref(txt), cnt)
    try:
        if rc != 0:
            raise RuntimeError("oracle_fold failed: %d" % rc)
        steps = parse_listing(C.string_at(txt).decode())
    finally:
        if txt.value:
            lib().oracle_free(txt)
    return (steps, list(cnt)) if counters else steps


def fold_batch(seqs, nb_mode=100, max_stack=1, max_branch=1000, min_hp=3, min_nrj=0.0, gc=3.0, au=2.0,
               gu=1.0, nthreads=None, want_text=True):
    """Thread pool over sequences.  -> (list of step lists or None, counters[n][CNT_N])."""
    import numpy as np
    n = len(seqs)
    offs = np.zeros(n + 1, dtype=np.int64)
    np.cumsum([len(s) for s in seqs], out=offs[1:])
    prm = Params(nb_mode, max_stack, max_branch, min_hp, 1, min_nrj, 37.0, gc, au, gu)
    texts = (C.c_void_p * n)() if want_text else None
    status = (C.c_int * n)()
    cnt = np.zeros((n, CNT_N), dtype=np.int64)
    nthreads = nthreads or os.cpu_count() or 1
    lib().oracle_fold_batch(C.byref(default_params()), "".join(seqs).encode(),
                            offs.ctypes.data_as(C.POINTER(C.c_int64)), n, C.byref(prm), nthreads,
                            texts, status, cnt.ctypes.data_as(C.POINTER(C.c_int64)))
    if any(status[i] for i in range(n)):
        raise RuntimeError("oracle_fold_batch: a sequence failed")
    out = None
    if want_text:
        out = []
        for i in range(n):
            out.append(parse_listing(C.string_at(texts[i]).decode()))
            lib().oracle_free(texts[i])
    return out, cnt


def fold_batch_check(seqs, raw, nb_mode=100, max_stack=1, max_branch=1000, min_hp=3, min_nrj=0.0, gc=3.0,
                     au=2.0, gu=1.0, traj=True, nthreads=None):
    """Fold `seqs` here and compare every beam (energies, dot-brackets, order) with `raw`, the arrays
    of a rafft_result (rafft_b200.rafft.RawResult), inside the C++ library -- full batches in the
    time of the oracle folds.  -> (indices of the sequences that differ, counters[n][CNT_N])."""
    import numpy as np
    n = len(seqs)
    assert raw.n_seq == n
    offs = np.zeros(n + 1, dtype=np.int64)
    np.cumsum([len(s) for s in seqs], out=offs[1:])
    prm = Params(nb_mode, max_stack, max_branch, min_hp, 1, min_nrj, 37.0, gc, au, gu)
    mism = np.zeros(max(n, 1), dtype=np.int32)
    cnt = np.zeros((n, CNT_N), dtype=np.int64)
    arrs = [np.ascontiguousarray(a) for a in (raw.status, raw.struct_off, raw.energy, raw.db_off, raw.db,
                                              raw.step_off, raw.entry_off, raw.entry_struct)]
    ptr = lambda a: C.c_void_p(a.ctypes.data)
    lib().oracle_fold_batch_check(C.byref(default_params()), "".join(seqs).encode(),
                                  offs.ctypes.data_as(C.POINTER(C.c_int64)), n, C.byref(prm),
                                  nthreads or os.cpu_count() or 1, 1 if traj else 0,
                                  *[ptr(a) for a in arrs], ptr(mism), ptr(cnt))
    return [int(i) for i in np.nonzero(mism[:n])[0]], cnt


def auto_cor(seq, pos, gc=3.0, au=2.0, gu=1.0):
    nv = len(pos)
    out = (C.c_double * (2 * nv - 1))()
    lib().oracle_auto_cor(seq.encode(), len(seq), (C.c_int * nv)(*pos), nv, gc, au, gu, out)
    return list(out)


def window_slide(seq, pos, lag, min_hp, gc=3.0, au=2.0, gu=1.0):
    nv = len(pos)
    res = (C.c_int * 3)()
    score = C.c_double()
    lib().oracle_window_slide(seq.encode(), len(seq), (C.c_int * nv)(*pos), nv, lag, min_hp, gc, au, gu, res,
                              C.byref(score))
    return res[0], res[1], res[2], score.value
