"""Test tiers: `-m "not gpu"` (oracle vs golden vectors, host logic, device-code emulation,
C-ABI symbols) runs on CPU; `-m gpu` (parity of the CUDA path with the oracle, through the
C-ABI) needs a B200.  Only this directory, smoke() and bench.py's CPU legs use oracle/."""
import gzip
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def oracle():
    import pyoracle
    pyoracle.lib()
    return pyoracle


@pytest.fixture(scope="session")
def native_lib_path():
    "the product library, built in-tree by nvcc (cross-compiles without a GPU)"
    from rafft_b200 import build
    return build.build()


@pytest.fixture(scope="session")
def ctx(native_lib_path):
    from rafft_b200 import rafft as R
    return R.get_context(0)


@pytest.fixture(scope="session")
def ref_folds():
    "trajectories of the UNMODIFIED reference Python (tools/make_golden.py)"
    with gzip.open(os.path.join(GOLDEN, "ref_folds.json.gz")) as fh:
        cases = json.load(fh)
    for c in cases:
        c["steps"] = [[(db, int(e)) for db, e in step] for step in c["steps"]]
    return cases


@pytest.fixture(scope="session")
def kernel_kat():
    with open(os.path.join(GOLDEN, "kernel_kat.json")) as fh:
        return json.load(fh)


@pytest.fixture(scope="session")
def energy_kat():
    import csv
    import io
    with gzip.open(os.path.join(GOLDEN, "energy_kat.csv.gz")) as fh:
        rows = list(csv.reader(io.StringIO(fh.read().decode())))
    return [(s, d, int(e)) for s, d, e in rows]


def parse_rafft_text(path):
    "example/rafft.out format -> (sequence, steps of (db, energy text))"
    lines = [l.rstrip("\n") for l in open(path)]
    start = 0
    seq = None
    if not lines[0].startswith("#"):
        seq, start = lines[0].strip(), 1
    steps = []
    for l in lines[start:]:
        if l.startswith("#"):
            steps.append([])
        elif l.strip():
            parts = l.split()
            if len(parts) == 1:  # older format repeats the sequence inside every block
                seq = parts[0]
                continue
            steps[-1].append((parts[0], parts[1]))
    return seq, steps


@pytest.fixture(scope="session")
def emu():
    "host emulation of the device code (tests/emu/emu_fold.cpp), test-only"
    import ctypes as C
    from rafft_b200._native import CNT_N, EnergyParams, Params
    bdir = os.path.join(ROOT, "tests", "emu", "_build")
    os.makedirs(bdir, exist_ok=True)
    so = os.path.join(bdir, "libemu.so")
    srcs = [os.path.join(ROOT, "tests", "emu", "emu_fold.cpp"),
            os.path.join(ROOT, "rafft_b200", "csrc", "fold_core.cuh"),
            os.path.join(ROOT, "rafft_b200", "csrc", "host_common.h")]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        r = subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", so, srcs[0]],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
    L = C.CDLL(so)
    L.emu_fold.argtypes = [C.POINTER(EnergyParams), C.c_char_p, C.c_int, C.POINTER(Params), C.c_int, C.c_int,
                           C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.POINTER(C.c_int32)]
    L.emu_fold.restype = C.c_int
    L.emu_free.argtypes = [C.c_void_p]
    L.emu_set_queue_cap.argtypes = [C.c_int]

    def fold(seq, reverse=0, nthreads=32, scale=1, queue_cap=0, **kw):
        import pyoracle
        L.emu_set_queue_cap(queue_cap)
        p = dict(nb_mode=100, max_stack=1, max_branch=1000, min_hp=3, min_nrj=0.0, gc=3.0, au=2.0, gu=1.0)
        p.update(kw)
        prm = Params(p["nb_mode"], p["max_stack"], p["max_branch"], p["min_hp"], 1, p["min_nrj"], 37.0,
                     p["gc"], p["au"], p["gu"])
        txt = C.c_void_p()
        cnt = (C.c_int64 * CNT_N)()
        hw = (C.c_int32 * 8)()
        rc = L.emu_fold(None, seq.encode(), len(seq), C.byref(prm), reverse, nthreads, scale, C.byref(txt),
                        cnt, hw)
        steps = pyoracle.parse_listing(C.string_at(txt).decode()) if txt.value else None
        if txt.value:
            L.emu_free(txt)
        return rc, steps, list(cnt), list(hw)
    return fold
