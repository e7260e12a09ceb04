"""Kernel-level parity (GPU) of the 16-bit activation locations: LayerNorm on int16 rows, the GEMM's
RESIDUAL / REQUANT epilogues with int16 outputs and identities, the stand-alone QuantAct over
int32 / int16 / int8 codes.  Bar: bit-exact against the numpy oracle (which is pinned against the
reference at 16 bit by tests/golden/vit_*_w16.npz)."""
import numpy as np
import pytest
import torch

from ivit_b200 import _cabi, hostmath as hm, ops
from oracle import ivit_oracle as O

pytestmark = pytest.mark.gpu
f32 = np.float32
DEV = "cuda"


def _cmp(got, exp, name):
    got = np.asarray(got).astype(np.int64)
    exp = np.asarray(exp).astype(np.int64)
    assert got.shape == exp.shape, (name, got.shape, exp.shape)
    bad = np.argwhere(got != exp)
    if len(bad):
        first = [(tuple(int(i) for i in b), int(got[tuple(b)]), int(exp[tuple(b)])) for b in bad[:8]]
        pytest.fail(f"{name}: {len(bad)}/{got.size} mismatches; first (idx, got, exp): {first}")


def _t(a, dtype):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dtype).to(DEV)


def _ln16_case(kind, C, rows, seed, s_in, spread, neg=False, use_int_sqrt=False, shift=0.0):
    rng = np.random.default_rng(seed)
    q = np.clip(np.rint(rng.standard_normal((1, rows, C)) * spread), -32768, 32767).astype(np.int64)
    q[0, 0, :] = 32767                      # saturated row: zero variance
    q[0, 1, ::2] = -32768                   # |y| > 46340: the int32 square wraps (hazard H9)
    q[0, 1, 1::2] = 32767
    for r in range(2, rows, 3):             # exact mean ties k + 1/2
        d = int(rng.integers(-3, 4)) * C + C // 2 - int(q[0, r].sum())
        idx = rng.integers(0, C, size=abs(d))
        np.add.at(q[0, r], idx, np.sign(d))
        q[0, r] = np.clip(q[0, r], -32768, 32767)
    s_in = f32(s_in)
    w = (rng.random(C) + 0.5).astype(f32)
    if neg:
        w[rng.random(C) < 0.1] *= -1
    b = (rng.standard_normal(C) * 0.1).astype(f32)
    x = (q.astype(f32) * s_in).astype(f32)
    if kind == 0:
        y, sf, _ = O.ivit_layernorm(x, np.array([s_in]), w, b)
    else:
        y, sf, _ = O.ibert_layernorm(x, np.array([s_in]), w, b, shift=shift, use_int_sqrt=use_int_sqrt)
    x_max = np.array([np.abs(y[np.isfinite(y)]).max() * 0.7], dtype=f32)
    _, s_out, codes = O.quant_act(y, 8, -x_max, x_max, pre_scale=sf)
    p = hm.layernorm_params(w, b, C)
    mult = hm.dyadic_mult(p["sf"], s_out)
    got = ops.layernorm_requant(kind, _t(q[0], torch.int16), float(s_in), _t(p["bias_int"], torch.float32),
                                _t(p["sf"], torch.float32), _t(mult, torch.float64), -128, 127, shift=shift,
                                use_int_sqrt=use_int_sqrt).cpu().numpy()
    ok = np.isfinite(y[0]).all(axis=1)      # rows whose reference output is inf/nan (zero variance) carry no code
    _cmp(got[ok], codes[0][ok], f"layernorm16 kind={kind} C={C}")
    return int(ok.sum())


@pytest.mark.parametrize("C", [96, 128, 192, 256, 384, 512, 768, 1024, 1536, 2048])
def test_layernorm16_ivit(C):
    assert _ln16_case(0, C, 203, seed=C, s_in=1.37e-4, spread=9000.0) >= 200
    _ln16_case(0, C, 64, seed=C + 1, s_in=3.1e-5, spread=20000.0, neg=True)


@pytest.mark.parametrize("C", [192, 256, 384, 512, 768])
def test_layernorm16_ivit_both_kernel_families(C):
    """The restructured kernels (x/s as two fused operations, taken for scales ivit_xeff_fma_ok accepts) and the division
    kernels (ivit_layernorm_set_variant(1)) against the same oracle rows."""
    assert _cabi.load().ivit_xeff_fma_ok(float(f32(1.37e-4)), 16) == 1
    try:
        for variant in (0, 1):
            ops.set_layernorm_variant(variant)
            _ln16_case(0, C, 203, seed=C + 7, s_in=1.37e-4, spread=9000.0)
            _ln16_case(0, C, 37, seed=C + 8, s_in=2.9e-5, spread=25000.0, neg=True)
    finally:
        ops.set_layernorm_variant(0)


@pytest.mark.parametrize("C", [192, 384, 768])
def test_layernorm16_ibert(C):
    try:
        for variant in (0, 1):              # restructured kernel / division kernel
            ops.set_layernorm_variant(variant)
            assert _ln16_case(1, C, 203, seed=C, s_in=1.37e-4, spread=9000.0, shift=5.0) >= 200
            _ln16_case(1, C, 64, seed=C + 1, s_in=3.1e-5, spread=3000.0, neg=True, shift=3.0)
            _ln16_case(1, C, 64, seed=C + 2, s_in=3.1e-5, spread=3000.0, use_int_sqrt=True, shift=3.0)
    finally:
        ops.set_layernorm_variant(0)


def test_layernorm16_strided_class_rows():
    """The final norm reads the class-token rows only: row stride N*C elements."""
    rng = np.random.default_rng(3)
    B, N, C = 5, 7, 192
    q = rng.integers(-20000, 20000, size=(B, N, C)).astype(np.int16)
    s_in = f32(2e-4)
    w = (rng.random(C) + 0.5).astype(f32)
    b = (rng.standard_normal(C) * 0.1).astype(f32)
    y, sf, _ = O.ivit_layernorm((q[:, :1].astype(f32) * s_in).astype(f32), np.array([s_in]), w, b)
    x_max = np.array([np.abs(y).max() * 0.8], dtype=f32)
    _, s_out, codes = O.quant_act(y, 8, -x_max, x_max, pre_scale=sf)
    p = hm.layernorm_params(w, b, C)
    mult = hm.dyadic_mult(p["sf"], s_out)
    xt = _t(q, torch.int16)
    got = ops.layernorm_requant(0, xt.view(B, N * C)[:, :C], float(s_in), _t(p["bias_int"], torch.float32),
                                _t(p["sf"], torch.float32), _t(mult, torch.float64), -128, 127)
    _cmp(got.cpu().numpy(), codes[:, 0], "layernorm16 class rows")


def _gemm16_case(M, N, K, seed, epilogue, bits1, bits2, id_bits, group=0):
    """QuantLinear -> QuantAct(bits1) [-> QuantAct(bits2) with an identity of id_bits]."""
    rng = np.random.default_rng(seed)
    A = rng.integers(-128, 128, size=(M, K)).astype(np.int8)
    W = rng.integers(-127, 128, size=(N, K)).astype(np.int8)
    bias = rng.integers(-20000, 20000, size=N).astype(np.int32)
    acc = (A.astype(np.float64) @ W.astype(np.float64).T + bias).astype(np.int64)
    s_in = (rng.random(N) * 2e-5 + 1e-5).astype(f32)
    y = (acc.astype(f32) * s_in).astype(f32)
    x_max = np.array([np.abs(y).max() * 0.6], dtype=f32)
    _, s1, c1 = O.quant_act(y.reshape(1, M, N), bits1, -x_max, x_max, pre_scale=s_in)
    m1 = hm.dyadic_mult(s_in, s1)
    lo1, hi1 = hm.qrange(bits1)
    kw = dict(bias=_t(bias, torch.int32), mult=_t(m1, torch.float64), s_in=_t(s_in, torch.float32), qmin=lo1, qmax=hi1,
              force_slow=bool(np.abs(m1).max() >= 256.0))
    exp, out_bits = c1[0], bits1
    if epilogue == _cabi.EPI_RESIDUAL:
        lim = 2 ** (id_bits - 1)
        rows_id = M if group == 0 else group + 1
        idc = rng.integers(-lim, lim, size=(rows_id, N)).astype(np.int16 if id_bits > 8 else np.int8)
        s_id = f32(0.021 * 128 / lim)
        x1 = (c1 * s1).astype(f32)
        if group == 0:
            ident = (idc.astype(f32) * s_id).astype(f32).reshape(1, M, N)
        else:
            ident = np.tile((idc[1:].astype(f32) * s_id).astype(f32), (M // group, 1)).reshape(1, M, N)
        x_max2 = np.array([np.abs(x1).max() * 0.9], dtype=f32)
        _, s2, c2 = O.quant_act(x1, bits2, -x_max2, x_max2, pre_scale=s1, identity=ident, identity_scale=np.array([s_id]))
        lo2, hi2 = hm.qrange(bits2)
        kw.update(mult2=float(hm.dyadic_mult(s1, s2).reshape(-1)[0]),
                  mult_id=float(hm.dyadic_mult(np.array([s_id]), s2).reshape(-1)[0]), qmin2=lo2, qmax2=hi2,
                  identity=_t(idc, torch.int16 if id_bits > 8 else torch.int8))
        exp, out_bits = c2[0], bits2
    odt = torch.int16 if out_bits > 8 else torch.int8
    out_rows = None
    if group:
        kw.update(group=group, out_group_stride=group + 1, out_row_off=1, id_broadcast=True)
        out_rows = (M // group) * (group + 1)
    for simple in (True, False):
        out = torch.full((out_rows or M, N), 99, dtype=odt, device=DEV)
        got = ops.gemm(_t(A, torch.int8), _t(W, torch.int8), epilogue, out=out, simple=simple, **kw).cpu().numpy()
        if group:
            got = got.reshape(M // group, group + 1, N)
            assert (got[:, 0] == 99).all()
            got = got[:, 1:].reshape(M, N)
        _cmp(got, exp, f"gemm16 {'simple' if simple else 'tcgen05'} M={M} N={N} K={K} bits=({bits1},{bits2},{id_bits})")


@pytest.mark.parametrize("shape", [(300, 384, 384), (197 * 2, 192, 768), (130, 129, 64), (1000, 384, 1536)])
def test_gemm16_residual_stream(shape):
    """attn.proj / mlp.fc2 with attention_out_bw = norm2_in_bw = 16 and an int16 identity."""
    _gemm16_case(*shape, seed=1, epilogue=_cabi.EPI_RESIDUAL, bits1=16, bits2=16, id_bits=16)


@pytest.mark.parametrize("shape", [(30000, 384, 384), (20011, 96, 96), (5000, 96, 384), (4099, 192, 768), (777, 100, 64)])
def test_gemm16_residual_stream_many_tiles(shape):
    """More tiles than SMs, ragged last token tile, Swin's narrow widths (N = 96: a partial channel block)."""
    _gemm16_case(*shape, seed=11, epilogue=_cabi.EPI_RESIDUAL, bits1=16, bits2=16, id_bits=16)
    _gemm16_case(*shape, seed=12, epilogue=_cabi.EPI_RESIDUAL, bits1=8, bits2=16, id_bits=16)


def test_gemm16_mixed_widths():
    _gemm16_case(300, 384, 384, seed=2, epilogue=_cabi.EPI_RESIDUAL, bits1=16, bits2=8, id_bits=8)
    _gemm16_case(300, 384, 384, seed=3, epilogue=_cabi.EPI_RESIDUAL, bits1=8, bits2=16, id_bits=16)
    _gemm16_case(300, 384, 384, seed=4, epilogue=_cabi.EPI_RESIDUAL, bits1=8, bits2=16, id_bits=8)
    _gemm16_case(300, 384, 384, seed=5, epilogue=_cabi.EPI_RESIDUAL, bits1=16, bits2=8, id_bits=16)
    _gemm16_case(300, 200, 384, seed=6, epilogue=_cabi.EPI_REQUANT, bits1=16, bits2=16, id_bits=16)
    _gemm16_case(300, 384, 384, seed=7, epilogue=_cabi.EPI_RESIDUAL, bits1=4, bits2=4, id_bits=8)


def test_gemm16_patch_embed_remap():
    _gemm16_case(196 * 3, 192, 768, seed=8, epilogue=_cabi.EPI_RESIDUAL, bits1=16, bits2=16, id_bits=16, group=196)


@pytest.mark.parametrize("z_dt,id_dt,out_bits", [(np.int32, np.int16, 16), (np.int16, np.int16, 16), (np.int16, None, 8),
                                                 (np.int8, np.int16, 16), (np.int32, np.int8, 16)])
def test_requant_codes(z_dt, id_dt, out_bits):
    rng = np.random.default_rng(5)
    rows, C = 77, 96
    lim = {np.int32: 3_000_000, np.int16: 32768, np.int8: 128}[z_dt]
    z = rng.integers(-lim, lim, size=(rows, C)).astype(z_dt)
    s_in = (rng.random(C) * 2e-5 + 1e-5).astype(f32) if z_dt == np.int32 else np.array([f32(3e-4)])
    y = (z.astype(f32) * s_in).astype(f32)
    x_max = np.array([np.abs(y).max() * 0.7], dtype=f32)
    ident = s_id = idc = None
    if id_dt is not None:
        il = 32768 if id_dt == np.int16 else 128
        idc = rng.integers(-il, il, size=(rows, C)).astype(id_dt)
        s_id = np.array([f32(x_max[0] / il)])
        ident = (idc.astype(f32) * s_id).astype(f32).reshape(1, rows, C)
    _, s_out, exp = O.quant_act(y.reshape(1, rows, C), out_bits, -x_max, x_max, pre_scale=s_in, identity=ident,
                                identity_scale=s_id)
    lo, hi = hm.qrange(out_bits)
    got = ops.requant(_t(z, {np.int32: torch.int32, np.int16: torch.int16, np.int8: torch.int8}[z_dt]),
                      _t(s_in, torch.float32), _t(hm.dyadic_mult(s_in, s_out), torch.float64), lo, hi,
                      identity=None if idc is None else _t(idc, torch.int16 if id_dt == np.int16 else torch.int8),
                      mult_id=0.0 if idc is None else float(hm.dyadic_mult(s_id, s_out).reshape(-1)[0]),
                      out_bits=out_bits)
    _cmp(got.cpu().numpy(), exp[0], "requant_codes")
