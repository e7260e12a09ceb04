"""Test infrastructure: shape-only stand-ins for the C-ABI launches (zeros of the right shape and type, a log of
what was asked for), so that the HOST logic of the frozen module path -- which kernel a call pattern maps to, with
which geometry -- can be walked on a machine without a GPU.  Nothing is computed; parity lives in the `-m gpu` tests."""
import numpy as np
import torch

from ivit_b200 import _cabi, ops
from ivit_b200.quantization_utils import frozen as fz


def install(monkeypatch):
    calls = []
    i8 = lambda bits: torch.int8 if bits <= 8 else torch.int16

    def quantize(x, inv, bits=8):
        calls.append("quantize")
        return torch.zeros(x.shape, dtype=i8(bits))

    def quantize_patches(x, inv, P, out=None):
        calls.append("quantize_patches")
        B, C, H, W = x.shape
        return torch.zeros((B * (H // P) * (W // P), C * P * P), dtype=torch.int8)

    def gemm(A, W, epi, *, out_bits=8, **k):
        calls.append(f"gemm:{epi}")
        if epi == _cabi.EPI_LOGITS:
            return torch.zeros((A.shape[0], W.shape[0]), dtype=torch.float32)
        return torch.zeros((A.shape[0], W.shape[0]), dtype=torch.int32 if epi == _cabi.EPI_ACC32 else i8(out_bits))

    def requant(z, s_in, mult, qmin, qmax, identity=None, mult_id=0.0, out=None, out_bits=8):
        calls.append(f"requant:{z.dtype}:{None if identity is None else identity.dtype}:{out_bits}")
        assert z.is_contiguous() and (identity is None or (identity.is_contiguous() and z.numel() % identity.numel() == 0))
        return torch.zeros(z.shape, dtype=i8(out_bits))

    def layernorm_requant(kind, x, s_in, bias_int, sf, mult, qmin, qmax, shift=0.0, use_int_sqrt=False, out=None):
        calls.append(f"ln:{x.dtype}")
        return torch.zeros(x.shape, dtype=torch.int8)

    def layernorm_requant_strided(kind, x, tokens, s_in, bias_int, sf, mult, qmin, qmax, shift=0.0, use_int_sqrt=False,
                                  out=None):
        calls.append(f"ln_strided:{x.dtype}:{tokens}")
        assert x.dim() == 2 and x.shape[0] % tokens == 0
        return torch.zeros(x.shape, dtype=torch.int8)

    def gelu_lut_build(kind, s_in, output_bit, mult, qmin, qmax, device):
        return torch.zeros((256, 256), dtype=torch.int8), float(np.float32(s_in) * np.float32(1 / 2 ** (output_bit - 1)))

    def gelu_lut_apply(x, lut, out=None):
        calls.append("gelu")
        return torch.zeros_like(x)

    def softmax_table_build(params, device, collapse=True, host_table=None):
        return torch.zeros((256, 256))

    def attention(qkv, B, N, H, hd, params, **k):
        calls.append(f"attention:{B}:{N}:{H}")
        return torch.zeros((B, N, H * hd), dtype=torch.int8)

    def window_attention(qkv, B, res, ws, shift, nH, hd, params, out=None, want_probs=False):
        calls.append(f"window_attention:B={B}:res={res}:ws={ws}:shift={shift}:nH={nH}:windowed={params.windowed}")
        assert qkv.numel() == B * res * res * 3 * nH * hd and params.bias_codes
        return torch.zeros((B, res * res, nH * hd), dtype=torch.int8)

    def avgpool_requant(x, B, L, C, s_in, mult, qmin, qmax, out=None):
        calls.append(f"avgpool:{L}:{C}")
        return torch.zeros((B, C), dtype=torch.int8)

    for name, fn in dict(quantize=quantize, quantize_patches=quantize_patches, gemm=gemm, requant=requant,
                         layernorm_requant=layernorm_requant, layernorm_requant_strided=layernorm_requant_strided,
                         gelu_lut_build=gelu_lut_build, gelu_lut_apply=gelu_lut_apply,
                         softmax_table_build=softmax_table_build, attention=attention, window_attention=window_attention,
                         avgpool_requant=avgpool_requant).items():
        monkeypatch.setattr(ops, name, fn)
    monkeypatch.setattr(fz, "require_cuda", lambda t, what: None)
    return calls
