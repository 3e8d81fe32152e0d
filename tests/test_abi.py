"""The C-ABI library builds for sm_100a, loads, and exports every symbol include/rafft_b200.h
declares; without a GPU it fails loudly instead of falling back.  No compute calls here."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "rafft_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(rafft_[a-z_0-9]+)\s*\(", txt)))


def test_header_symbols_exported(native_lib_path):
    from rafft_b200 import _native
    syms = header_symbols()
    assert sorted(_native.EXPORTS) == syms
    L = C.CDLL(native_lib_path)
    for s in syms:
        assert hasattr(L, s), s


def test_version_and_default_tables(native_lib_path):
    from rafft_b200 import _native
    assert b"sm_100a" in _native.lib().rafft_version()
    ep = _native.default_energy_params()
    assert ep.terminal_au == 50 and ep.ml_closing == 930 and ep.ninio == 60
    assert ep.stack[0] == -240 and ep.hairpin[3] == 540 and ep.n_tetra == 16


def test_struct_layouts_match_header(native_lib_path, tmp_path):
    "ctypes mirrors of the ABI structs have the sizes the C compiler gives them"
    from rafft_b200 import _native
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "rafft_b200.h"\nint main(){printf("%zu %zu %zu\\n",'
                   'sizeof(rafft_energy_params),sizeof(rafft_params),sizeof(rafft_result));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    a, b, c = map(int, subprocess.check_output([str(exe)]).split())
    assert (a, b, c) == (C.sizeof(_native.EnergyParams), C.sizeof(_native.Params), C.sizeof(_native.Result))


def test_cuda_sass_is_sm100a(native_lib_path):
    out = subprocess.run(["cuobjdump", "-lelf", native_lib_path], capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_no_cpu_fallback_without_gpu(native_lib_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from rafft_b200 import _native, fold
    with pytest.raises(_native.NativeError) as ei:
        fold("GGGAAACCC")
    assert ei.value.code == _native.RAFFT_ERR_CUDA and "no CPU fallback" in str(ei.value)


def test_product_does_not_reference_oracle():
    "nothing in the shipped package imports, links or names oracle/"
    for pkg in ("rafft_b200", "rafft", "bin", "include"):
        for dp, _, fs in os.walk(os.path.join(ROOT, pkg)):
            for f in fs:
                if f.endswith((".py", ".cu", ".cuh", ".h")) or pkg == "bin":
                    txt = open(os.path.join(dp, f), errors="replace").read()
                    assert "pyoracle" not in txt and "liboracle" not in txt and "oracle_" not in txt, f
