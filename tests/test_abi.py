"""The C-ABI seam without a GPU: the in-tree shared library loads with plain `ctypes`, exports every entry point that
include/dphubert_b200.h declares, the ctypes signature table covers the same set, and calls fail loudly (no compute
is attempted here)."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "dphubert_b200.h")
LIB = os.path.join(ROOT, "dphubert_b200", "libdphubert_b200.so")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)  # drop comments
    return sorted(set(re.findall(r"\b(dph_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(LIB):
        from dphubert_b200 import _build
        _build.build()
    return ctypes.CDLL(LIB)


def test_header_declares_the_path():
    syms = declared_symbols()
    assert len(syms) >= 45
    for must in ("dph_gemm_bf16", "dph_attn_fwd", "dph_attn_bwd", "dph_conv0_fwd", "dph_conv0_bwd", "dph_layernorm_fwd",
                 "dph_hardconcrete_fwd", "dph_num_params", "dph_distill_loss_fwd", "dph_adamw", "dph_last_error"):
        assert must in syms, must


def test_library_exports_every_declared_symbol(lib):
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, "declared in include/dphubert_b200.h but not exported: %s" % missing


def test_no_undeclared_entry_points():
    out = subprocess.run(["nm", "-D", "--defined-only", LIB], capture_output=True, text=True, check=True).stdout
    exported = sorted(set(re.findall(r" T (dph_[a-z0-9_]+)$", out, flags=re.M)))
    extra = [s for s in exported if s not in declared_symbols()]
    assert not extra, "exported but not declared in the header: %s" % extra


def test_ctypes_signature_table_matches_header():
    from dphubert_b200 import _sigs
    table = set(_sigs.SIGS)
    declared = set(declared_symbols())
    # entry points that take a struct pointer or no tensors are bound by hand in _lib.py
    by_hand = {"dph_gemm_bf16", "dph_gemm_f32", "dph_last_error", "dph_launch_count", "dph_gemm_debug_timeline"}
    assert declared - table - by_hand == set(), sorted(declared - table - by_hand)
    assert table - declared == set(), sorted(table - declared)


def test_error_reporting_without_a_device(lib):
    lib.dph_last_error.restype = ctypes.c_char_p
    lib.dph_gemm_bf16.restype = ctypes.c_int
    rc = lib.dph_gemm_bf16(None, None)  # null descriptor: argument error, no launch
    assert rc < 0
    assert b"null" in lib.dph_last_error().lower()
