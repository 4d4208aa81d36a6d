"""CPU: the C-ABI shared library builds for sm_100a, loads, and exports exactly the symbols that
include/ggpm_b200.h declares; the ctypes signature table covers all of them.  No compute is called."""
import ctypes
import os
import re

import pytest


def header_symbols(root_dir):
    text = open(os.path.join(root_dir, "include", "ggpm_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ggpm_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def lib_path():
    from ggpm_b200 import build
    return build.build()


def test_library_exports_every_declared_symbol(root_dir, lib_path):
    lib = ctypes.CDLL(lib_path)
    syms = header_symbols(root_dir)
    assert len(syms) >= 15
    for s in syms:
        assert hasattr(lib, s), "missing export " + s


def test_ctypes_table_matches_header(root_dir, lib_path):
    from ggpm_b200 import _lib
    assert sorted(_lib.SIGNATURES) == header_symbols(root_dir)
    lib = _lib._load()
    assert lib.ggpm_abi_version() == 1
    assert lib.ggpm_launch_count() == 0 or lib.ggpm_launch_count() > 0
    assert lib.ggpm_get_gemm_engine() in (0, 1, 2, 3)


def test_argument_validation_needs_no_gpu(lib_path):
    """errors that are detected before any CUDA call: status != 0 and a message"""
    from ggpm_b200 import _lib
    lib = _lib._load()
    assert lib.ggpm_set_gemm_engine(99) != 0
    assert b"engine" in lib.ggpm_last_error()
    assert lib.ggpm_gather_sum(None, 4, None, None, None, 4, 1, 4, 0, None) != 0
    assert b"null" in lib.ggpm_last_error()
    assert lib.ggpm_lstm_fwd(5, 10, 33, 2, *([None] * 15), 0, 0, None) != 0
    assert b"multiple of 32" in lib.ggpm_last_error()


def test_sass_is_sm100a(lib_path):
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", lib_path], capture_output=True, text=True).stdout
    assert "sm_100a" in out
