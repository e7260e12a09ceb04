"""ivit_gelu_i8_noalloc (GPU): the allocation-free GELU entry point against the existing two-call form
(ivit_gelu_lut_build + ivit_gelu_lut_apply_i8), through the C ABI, bit-exact (the tolerance of every integer
test here), on two calls in a row that share the library's static scratch.

Collected last on purpose (file name).
"""
import ctypes as C

import numpy as np
import pytest
import torch

from ivit_b200 import _cabi, hostmath as hm

pytestmark = pytest.mark.gpu
DEV = "cuda"
MAX_ROWS = 65536      # IVIT_GELU_NOALLOC_MAX_ROWS (include/ivit_b200.h)


def _two_call(lib, kind, x, s_in, mult, st):
    rows, Cdim = x.shape
    lut = torch.empty((256, 256), dtype=torch.int8, device=DEV)
    out = torch.empty_like(x)
    s_out = C.c_float(0)
    _cabi.check(lib.ivit_gelu_lut_build(kind, _cabi.ptr(lut), C.c_float(s_in), 8, C.c_double(mult), -128, 127,
                                        C.byref(s_out), st), "ivit_gelu_lut_build")
    _cabi.check(lib.ivit_gelu_lut_apply_i8(_cabi.ptr(x), _cabi.ptr(out), _cabi.ptr(lut), rows, Cdim, st),
                "ivit_gelu_lut_apply_i8")
    return out, s_out.value


def _no_alloc(lib, kind, x, out, s_in, mult, st):
    rows, Cdim = x.shape
    s_out = C.c_float(0)
    _cabi.check(lib.ivit_gelu_i8_noalloc(kind, _cabi.ptr(x), _cabi.ptr(out), rows, Cdim, C.c_float(s_in), 8,
                                         C.c_double(mult), -128, 127, C.byref(s_out), st), "ivit_gelu_i8_noalloc")
    return s_out.value


def test_gelu_noalloc_matches_two_call_form_twice(lib):
    g = torch.Generator().manual_seed(7)
    st = _cabi.stream()
    # two different layers back to back: kind, scales, widths and row counts all change, so the second call
    # finds a stale table and stale row maxima in the shared scratch
    calls = [(0, 0.0431, 0.0197, 1031, 1536), (1, 0.0127, 0.0093, 397, 3072), (0, 0.0431, 0.0197, 1031, 1536)]
    for kind, s_in, s_next, rows, Cdim in calls:
        x = torch.randint(-128, 128, (rows, Cdim), generator=g, dtype=torch.int8)
        x[::7] = torch.clamp(x[::7], max=-3)        # rows whose maximum is negative
        x = x.to(DEV)
        _, s_gelu = _two_call(lib, kind, x, s_in, 1.0, st)           # the module's output scale, for mlp.qact1's ratio
        mult = float(np.asarray(hm.dyadic_mult(np.float32(s_gelu), np.float32(s_next))).reshape(-1)[0])
        want, s_gelu = _two_call(lib, kind, x, s_in, mult, st)
        out = torch.full_like(x, 55)
        before = torch.cuda.memory_allocated()
        s_gelu2 = _no_alloc(lib, kind, x, out, s_in, mult, st)
        assert torch.cuda.memory_allocated() == before
        torch.cuda.synchronize()
        assert s_gelu2 == s_gelu
        assert torch.equal(out, want), f"kind {kind}: {(out != want).sum().item()} of {out.numel()} codes differ"
        assert out.unique().numel() > 8


def test_gelu_noalloc_rejects_oversize_and_bad_shapes(lib):
    st = _cabi.stream()
    x = torch.zeros((8, 64), dtype=torch.int8, device=DEV)
    rc = lib.ivit_gelu_i8_noalloc(0, _cabi.ptr(x), _cabi.ptr(x), MAX_ROWS + 1, 64, C.c_float(0.05), 8,
                                  C.c_double(0.5), -128, 127, None, st)
    assert rc == -2 and b"IVIT_GELU_NOALLOC_MAX_ROWS" in lib.ivit_last_error()
    rc = lib.ivit_gelu_i8_noalloc(0, _cabi.ptr(x), _cabi.ptr(x), 8, 60, C.c_float(0.05), 8, C.c_double(0.5), -128, 127,
                                  None, st)
    assert rc == -2
    rc = lib.ivit_gelu_i8_noalloc(2, _cabi.ptr(x), _cabi.ptr(x), 8, 64, C.c_float(0.05), 8, C.c_double(0.5), -128, 127,
                                  None, st)
    assert rc == -1
    # zero rows: a no-op that still succeeds
    assert lib.ivit_gelu_i8_noalloc(0, _cabi.ptr(x), _cabi.ptr(x), 0, 64, C.c_float(0.05), 8, C.c_double(0.5), -128,
                                    127, None, st) == 0
    torch.cuda.synchronize()
