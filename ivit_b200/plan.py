"""Frozen static plan of the ViT/DeiT integer forward (the hot path BASELINE.json names).

After ``freeze_model`` (model_utils.py:5) every scale, dyadic multiplier, integer weight, integer
bias, LayerNorm ``bias_int`` and GELU / softmax constant is a function of (weights,
``x_min``/``x_max`` buffers) only -- yet the reference recomputes all of it on every forward
(quant_modules.py:202-220, :371-372; quant_utils.py:151-175 with a device->host sync per
QuantAct).  ``FrozenVitPlan`` computes it ONCE on the host (``hostmath``: the reference's exact
float32/float64 operation order), uploads int8 weights / int32 biases / float64 multipliers, and
replays ``VisionTransformer.forward`` (vit_quant.py:285-312) as ~100 launches of the sm_100a
kernels behind ``include/ivit_b200.h``; the launch sequence can be captured into a CUDA graph.

Kernel fusion map (reference modules -> one launch):
  qact_input + im2col                                   -> ivit_quantize_patches_f32_i8
  patch_embed.proj + patch_embed.qact + qact1(+pos)     -> ivit_gemm_i8 (RESIDUAL, row remap)
  norm1 + Block.qact1 | norm2 + qact3 | norm + qact2    -> ivit_layernorm_requant_i8
  attn.qkv + attn.qact1                                 -> ivit_gemm_i8 (REQUANT)
  matmul_1 * scale + qact_attn1 + int_softmax + matmul_2 + attn.qact2 -> ivit_attention_i8
  attn.proj + attn.qact3 + Block.qact2(identity)        -> ivit_gemm_i8 (RESIDUAL)
  mlp.fc1 + mlp.qact_gelu                               -> ivit_gemm_i8 (REQUANT)
  mlp.act (ShiftGELU | I-BERT GELU) + mlp.qact1         -> ivit_gelu_lut_apply_i8
  mlp.fc2 + mlp.qact2 + Block.qact4(identity)           -> ivit_gemm_i8 (RESIDUAL)
  head                                                  -> ivit_gemm_i8 (LOGITS)

There is no CPU path: every method that computes raises unless the tensors are on a CUDA device
and the in-tree library is present.
"""
from __future__ import annotations

import os

from dataclasses import dataclass, field

import numpy as np
import torch

from . import _cabi, hostmath as hm, ops
from ._cabi import EPI_LOGITS, EPI_REQUANT, EPI_RESIDUAL, AttentionParams
from .synthetic import VitConfig

f32 = np.float32


def _raw(v) -> np.ndarray:
    return v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)


def _np(v) -> np.ndarray:
    if isinstance(v, torch.Tensor):
        v = v.detach().cpu().numpy()
    return np.asarray(v, dtype=np.float32)


@dataclass
class _Linear:
    """QuantLinear (+ the QuantAct that follows it), frozen."""
    w: torch.Tensor            # int8 [N, K]
    bias: torch.Tensor         # int32 [N]
    b_scale: torch.Tensor      # float32 [N]  per-channel scale of acc+bias (s_w * s_x)
    mult: torch.Tensor | None  # float64 [N]  m*2^-e of b_scale / s_out
    qmin: int = -128
    qmax: int = 127
    force_slow: bool = False
    z_bound_ok: bool = False
    s_out: np.float32 = f32(0)
    # second requant + identity (RESIDUAL)
    mult2: float = 0.0
    mult_id: float = 0.0
    qmin2: int = -128
    qmax2: int = 127
    s_out2: np.float32 = f32(0)


@dataclass
class _Norm:
    kind: int
    bias_int: torch.Tensor
    sf: torch.Tensor
    mult: torch.Tensor
    s_out: np.float32
    shift: float = 0.0
    use_int_sqrt: bool = False
    qmin: int = -128
    qmax: int = 127


@dataclass
class _Block:
    norm1: _Norm = None
    qkv: _Linear = None
    attn: AttentionParams = None
    proj: _Linear = None
    norm2: _Norm = None
    fc1: _Linear = None
    gelu_lut: torch.Tensor = None
    fc2: _Linear = None
    s_in: np.float32 = f32(0)      # scale of the block input codes
    s_mid: np.float32 = f32(0)     # scale after Block.qact2
    s_out: np.float32 = f32(0)     # scale after Block.qact4
    scales: dict = field(default_factory=dict)


def _layer_kind(name: str) -> int:
    base = name.split("_")[0]
    if base == "ivit":
        return 0
    if base == "ibert":
        return 1
    if base == "ppoly":            # piecewise-polynomial GELU / softmax: per-layer tables from the stored coefficients
        return 2
    raise NotImplementedError(f"layer type {name!r}: this plan covers the ivit|ibert|ppoly registry entries")


def _device_guard(device):
    """Make the plan's device current for the duration of a build / launch sequence: the C ABI enqueues on the
    CURRENT device's stream and caches per-device launch attributes (a plan may be built for cuda:1 while cuda:0 is
    current).  No-op for the host-only plan."""
    import contextlib
    return torch.cuda.device(device) if torch.device(device).type == "cuda" else contextlib.nullcontext()


class GraphCaptureMixin:
    """Launch accounting and CUDA-graph capture shared by the frozen plans (`_enqueue(images)` enqueues one
    forward on torch's current stream)."""

    def _mark(self, kernel: str, ops_: float, bytes_: float):
        """One kernel launched.  With ``self.trace`` set (bench.py's per-kernel pass) a CUDA event is
        recorded after it, with the launch's ALGORITHMIC work: integer ops (1 MAC = 2) and the bytes
        of every operand read once + every result written once."""
        self.launches_per_forward += 1
        if self.trace is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self.trace.append((kernel, float(ops_), float(bytes_), ev))

    def set_uint8_input(self, mean, std):
        """Accept uint8 images [B,C,H,W] (0..255) as well: ToTensor + Normalize(mean, std) + the input QuantAct become
        one [C][256] code table (ops.input_code_lut; float32, torchvision's op order), applied by the input kernel."""
        self.input_lut = ops.dev(ops.input_code_lut(self.inv_s_input, mean, std), torch.int8, self.device)

    def capture(self, B: int, slot: int = 0, dtype=torch.float32) -> "GraphedForward":
        """Capture the launch sequence for batch B into a CUDA graph over a static input buffer.
        Graphs of different ``slot`` have their own input buffer but share the activation buffers
        (replay them on one stream)."""
        if (B, slot, dtype) in self._graphs:
            return self._graphs[(B, slot, dtype)]
        with _device_guard(self.device):
            return self._capture(B, slot, dtype)

    def _capture(self, B: int, slot: int, dtype) -> "GraphedForward":
        cfg = self.cfg
        static_in = torch.zeros((B, cfg.in_chans, cfg.img_size, cfg.img_size), dtype=dtype, device=self.device)
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):           # warm-up: cudaFuncSetAttribute etc. outside capture
            self._enqueue(static_in)
            self._enqueue(static_in)
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            out = self._enqueue(static_in)
        gf = GraphedForward(g, static_in, out, self.launches_per_forward)
        self._graphs[(B, slot, dtype)] = gf
        return gf



class FrozenVitPlan(GraphCaptureMixin):
    """Host-side constants + launch sequence of one frozen ViT/DeiT model.

    cfg: ``VitConfig``; state: reference ``state_dict`` names -> arrays (weights, biases,
    ``cls_token``, ``pos_embed``, and the calibrated ``*.x_min`` / ``*.x_max`` [/ ``*.shift``]).
    """

    def __init__(self, cfg: VitConfig, state: dict, device="cuda", host_only: bool = False):
        """host_only=True builds the host constants only (no kernels can run: used by the CPU tests
        that check the scale wiring); otherwise a CUDA device is mandatory."""
        self.cfg = cfg
        self.host_only = host_only
        self.device = torch.device("cpu" if host_only else device)
        if not host_only and self.device.type != "cuda":
            raise _cabi.IvitKernelError("FrozenVitPlan needs a CUDA device (there is no CPU fallback)")
        if len(cfg.bitwidths) != 8 or any(not 2 <= int(b) <= 16 for b in cfg.bitwidths):
            raise NotImplementedError("FrozenVitPlan: the 8 activation locations take 2..16 bits")
        # vit_quant.py:180-187: patch_embed, pos_encoding, block_input, attention_out, softmax, mlp_out,
        # norm2_in, att_block_out.  The residual stream (block_input / norm2_in / att_block_out) is stored
        # as int16 codes as soon as one of its three locations is wider than 8 bits, the positional
        # codes likewise; attention_out / mlp_out / patch_embed are in-register intermediates of a fused
        # epilogue and only change its clamp range; softmax > 8 bits selects the two-plane P.V kernel.
        (self.bw_patch, self.bw_pos, self.bw_block_in, self.bw_attn_out, self.bw_softmax, self.bw_mlp_out,
         self.bw_norm2_in, self.bw_block_out) = (int(b) for b in cfg.bitwidths)
        self.stream_dtype = torch.int16 if max(self.bw_block_in, self.bw_norm2_in, self.bw_block_out) > 8 else torch.int8
        self.pos_dtype = torch.int16 if self.bw_pos > 8 else torch.int8
        if not host_only:
            _cabi.load()
        # ppoly coefficient buffers (`*.fixed_bounds`, `*.fixed_coeffs`) are int32: keep them exact
        self.sd = {k: (_raw(v) if ".fixed_" in k else _np(v)) for k, v in state.items()
                   if not k.endswith(("weight_integer", "bias_integer"))}
        self.gelu_kind = _layer_kind(cfg.gelu_type)
        self.softmax_kind = _layer_kind(cfg.softmax_type)
        self.ln_kind = _layer_kind(cfg.layernorm_type)
        if self.ln_kind == 2:
            raise NotImplementedError("there is no ppoly LayerNorm (layer_selection.py:132-136)")
        from .quantization_utils.registry import parse_layer_name
        self.gelu_args = parse_layer_name(cfg.gelu_type)[1]
        self.softmax_args = parse_layer_name(cfg.softmax_type)[1]
        self.ln_int_sqrt = "use-int-sqrt_true" in cfg.layernorm_type
        self.C = cfg.embed_dim
        self.N = cfg.tokens
        self.H = cfg.num_heads
        self.hd = cfg.embed_dim // cfg.num_heads
        self._bufs = {}
        self._tables = []
        self._graphs = {}
        self.launches_per_forward = 0
        self.trace = None
        self.fuse_gelu = os.environ.get("IVIT_FUSE_GELU", "0") == "1"      # fc1 + GELU table in one launch (ivit_gemm_epilogue_t.gelu_lut); False: separate pass
        with _device_guard(self.device):        # constants, tables and launches on the plan's device, whatever is current
            self._build()

    # ------------------------------------------------------------------ host constants
    def _act_scale(self, name: str, bits: int = 8) -> np.ndarray:
        """quant_modules.py:368-372: the frozen QuantAct's scale from its x_min / x_max buffers."""
        return hm.sym_scale(bits, self.sd[name + ".x_min"], self.sd[name + ".x_max"]).reshape(1)

    def _dev(self, a, dtype):
        return ops.dev(a, dtype, self.device)

    def _linear(self, name: str, s_x: np.ndarray, s_out: np.ndarray | None, bits: int = 8) -> _Linear:
        p = hm.linear_params(self.sd[name + ".weight"], self.sd.get(name + ".bias"), s_x)
        mult = None
        slow = False
        if s_out is not None:
            m = hm.dyadic_mult(p["b_scale"], s_out)
            # fast epilogue: one fused fp64 FMA, valid while |z| < 2^22 and |z*mult| < 2^31
            slow = bool(np.abs(m).max() >= 256.0) or bool(np.abs(p["b_int_f32"] if p["b_int_f32"] is not None
                                                               else 0).max() >= 2 ** 30)
            mult = self._dev(m.reshape(-1), torch.float64)
        lo, hi = hm.qrange(bits)
        # weights-only bound on |acc + bias| for int8 inputs (|x| <= 128): below 2^22 the epilogue's
        # fast requant is exact for every possible activation and needs no per-element range check
        zb = np.abs(p["w_int_f32"]).sum(axis=1).max() * 128.0 + np.abs(p["b_int"].astype(np.float64)).max()
        self_zb = bool(zb < 2 ** 22) and not slow
        return _Linear(z_bound_ok=self_zb, w=self._dev(p["w_int"], torch.int8), bias=self._dev(p["b_int"], torch.int32),
                       b_scale=self._dev(p["b_scale"], torch.float32), mult=mult, qmin=lo, qmax=hi,
                       force_slow=slow, s_out=(s_out[0] if s_out is not None else f32(0)))

    def _residual(self, lin: _Linear, s_mid: np.ndarray, s_id: np.ndarray, s_out: np.ndarray, bits: int = 8):
        """Second fixedpoint_mul with the identity branch (quant_utils.py:232-245)."""
        lin.mult2 = float(hm.dyadic_mult(s_mid, s_out).reshape(-1)[0])
        lin.mult_id = float(hm.dyadic_mult(s_id, s_out).reshape(-1)[0])
        lin.qmin2, lin.qmax2 = hm.qrange(bits)
        lin.s_out2 = s_out[0]
        if max(abs(lin.mult2), abs(lin.mult_id)) >= 2 ** 20:
            lin.force_slow = True
        return lin

    def _norm(self, name: str, s_out: np.ndarray) -> _Norm:
        p = hm.layernorm_params(self.sd[name + ".weight"], self.sd[name + ".bias"], self.C)
        mult = hm.dyadic_mult(p["sf"], s_out).reshape(-1)
        shift = float(self.sd.get(name + ".shift", np.zeros(1, dtype=f32)).reshape(-1)[0])
        return _Norm(kind=self.ln_kind, bias_int=self._dev(p["bias_int"], torch.float32),
                     sf=self._dev(p["sf"], torch.float32), mult=self._dev(mult, torch.float64), s_out=s_out[0],
                     shift=shift, use_int_sqrt=self.ln_int_sqrt)

    def _attention(self, p: str, s_qkv: np.ndarray) -> tuple:
        """vit_quant.py:72-85: scales of matmul_1, *scale, qact_attn1, int_softmax, matmul_2, qact2."""
        a = AttentionParams()
        ob = self.bw_softmax
        a.kind, a.output_bit = self.softmax_kind, ob
        s_mm1 = (s_qkv * s_qkv).astype(f32)                        # quant_modules.py:409
        s_scaled = (s_mm1 * f32(self.hd ** -0.5)).astype(f32)      # vit_quant.py:75
        s_attn = self._act_scale(p + "qact_attn1")
        a.mult_attn = float(hm.dyadic_mult(s_scaled, s_attn).reshape(-1)[0])
        a.qmin_attn, a.qmax_attn = -128, 127
        a.s_attn = float(s_attn[0])
        host_table = None
        if self.softmax_kind == 0:
            s_sm = np.array([1 / 2 ** (ob - 1)], dtype=f32)        # ivit_modules.py:177
        elif self.softmax_kind == 2:
            from .quantization_utils import ppoly as pp
            s_sm = np.array([2 / 2 ** ob], dtype=f32)              # ppoly_modules.py:329
            host_table = pp.softmax_term_table(s_attn, self.sd[p + "int_softmax.fixed_bounds"],
                                               self.sd[p + "int_softmax.fixed_coeffs"],
                                               int(self.softmax_args.get("exp_bits", 16)))
        else:
            s_sm = np.array([2 / 2 ** ob], dtype=f32)              # ibert_modules.py:317
            s16 = self._act_scale(p + "int_softmax.act", 16)       # inner QuantAct(16), :262,:310
            exp_scale = ((f32(0.35815147) * (s_attn * s_attn)) / f32(2 ** 30)).astype(f32)
            a.mult_exp16 = float(hm.dyadic_mult(exp_scale, s16).reshape(-1)[0])
            a.s_exp16 = float(s16[0])
        s_mm2 = (s_sm * s_qkv).astype(f32)
        a.s_pv = float(s_mm2[0])
        s_out = self._act_scale(p + "qact2")
        a.mult_out = float(hm.dyadic_mult(s_mm2, s_out).reshape(-1)[0])
        a.qmin_out, a.qmax_out = -128, 127
        if not self.host_only:
            self._tables.append(ops.softmax_table_build(a, self.device, host_table=host_table))
        return a, s_out

    def _build(self):
        cfg, sd = self.cfg, self.sd
        C = self.C
        # ---- stem: qact_input -> patch_embed.proj -> patch_embed.qact -> cat(cls) -> qact1(+pos)
        self.s_input = self._act_scale("qact_input")
        self.inv_s_input = float(f32(1.0) / self.s_input[0])        # quant_utils.py:49 reciprocal
        s_pe = self._act_scale("patch_embed.qact", self.bw_patch)
        s_pos = self._act_scale("qact_pos", self.bw_pos)
        s_x = self._act_scale("qact1", self.bw_block_in)
        pe = self._linear("patch_embed.proj", self.s_input, s_pe, self.bw_patch)
        self.patch = self._residual(pe, s_pe, s_pos, s_x, self.bw_block_in)
        pos_codes = hm.quantize_tensor(sd["pos_embed"].reshape(self.N, C), self.bw_pos, s_pos)      # qact_pos
        self.pos_codes = self._dev(pos_codes, self.pos_dtype)
        # class-token row: the raw fp32 parameter enters qact1 un-quantised (vit_quant.py:290-296):
        # z = round(cls / s_pe) is NOT clamped to the patch_embed range.
        z_cls = hm.recover_codes(sd["cls_token"].reshape(1, C), s_pe)
        cls_row = hm.requant_host(z_cls, s_pe, s_x, self.bw_block_in, identity_codes=pos_codes[:1], s_id=s_pos)
        self.cls_row = self._dev(cls_row.reshape(C), self.stream_dtype)
        # ---- blocks
        self.blocks = []
        for i in range(cfg.depth):
            p = f"blocks.{i}."
            b = _Block(s_in=s_x[0])
            s_n1 = self._act_scale(p + "qact1")
            b.norm1 = self._norm(p + "norm1", s_n1)
            s_qkv = self._act_scale(p + "attn.qact1")
            b.qkv = self._linear(p + "attn.qkv", s_n1, s_qkv)
            b.attn, s_ao = self._attention(p + "attn.", s_qkv)
            s_q3 = self._act_scale(p + "attn.qact3", self.bw_attn_out)
            s_mid = self._act_scale(p + "qact2", self.bw_norm2_in)
            b.proj = self._residual(self._linear(p + "attn.proj", s_ao, s_q3, self.bw_attn_out), s_q3, s_x, s_mid,
                                    self.bw_norm2_in)
            s_n2 = self._act_scale(p + "qact3")
            b.norm2 = self._norm(p + "norm2", s_n2)
            s_g = self._act_scale(p + "mlp.qact_gelu")
            b.fc1 = self._linear(p + "mlp.fc1", s_n2, s_g)
            s_m1 = self._act_scale(p + "mlp.qact1")
            b.gelu_lut, b.scales["gelu_out"] = self._gelu_lut(s_g, s_m1, p + "mlp.act")
            s_m2 = self._act_scale(p + "mlp.qact2", self.bw_mlp_out)
            s_o = self._act_scale(p + "qact4", self.bw_block_out)
            b.fc2 = self._residual(self._linear(p + "mlp.fc2", s_m1, s_m2, self.bw_mlp_out), s_m2, s_mid, s_o,
                                   self.bw_block_out)
            b.s_mid, b.s_out = s_mid[0], s_o[0]
            b.scales.update(n1=s_n1[0], qkv=s_qkv[0], attn_out=s_ao[0], n2=s_n2[0], gelu_in=s_g[0], mlp1=s_m1[0])
            self.blocks.append(b)
            s_x = s_o
        # ---- final norm (class rows only) -> qact2 -> head
        s_f = self._act_scale("qact2")
        self.norm = self._norm("norm", s_f)
        self.head = self._linear("head", s_f, None)
        self.s_final = s_x[0]
        if not self.host_only:
            torch.cuda.synchronize(self.device)

    def _gelu_lut(self, s_in: np.ndarray, s_out: np.ndarray, name: str = ""):
        """mlp.act + mlp.qact1 tabulated over (row max code, code): built on device by the exact
        arithmetic (rowwise.cu gelu_lut_build_kernel).  The dyadic multiplier needs the GELU's own
        output scale first: ivit s/2^7 (ivit_modules.py:121-124), ibert s*s_erf/2 (:233)."""
        s = f32(s_in[0])
        if self.gelu_kind == 2:        # ppoly: the module's stored coefficients, tabulated on the host
            from .quantization_utils.ppolymods import gelu_lut_from_coeffs
            s_g = np.asarray(self.sd[name + ".fixed_scaling_factor_out"], dtype=f32).reshape(1)
            if self.host_only:
                return None, s_g[0]
            lut = gelu_lut_from_coeffs(s, self.sd[name + ".fixed_bounds"], self.sd[name + ".fixed_coeffs"], s_g,
                                       int(self.gelu_args.get("scale_bits", 22)), s_out, 8, self.device)
            return lut, s_g[0]
        if self.gelu_kind == 0:
            s_g = np.array([s * f32(1 / 2 ** 7)], dtype=f32)
        else:
            k = f32(s / f32(1.4142))
            s_erf = f32(f32(f32(k * k) * f32(-0.2888)) * f32(64.0))
            s_g = np.array([f32(f32(s * s_erf) / f32(2.0))], dtype=f32)
        mult = float(hm.dyadic_mult(s_g, s_out).reshape(-1)[0])
        if self.host_only:
            return None, s_g[0]
        lut, s_g_dev = ops.gelu_lut_build(self.gelu_kind, float(s), 8, mult, -128, 127, self.device)
        if f32(s_g_dev) != s_g[0]:
            raise _cabi.IvitKernelError(f"GELU output scale mismatch host {s_g[0]!r} vs library {s_g_dev!r}")
        return lut, s_g[0]

    # ------------------------------------------------------------------ device buffers
    def _buffers(self, B: int) -> dict:
        """Activation buffers for batch B (allocated once, reused by every forward / graph replay).
        int8 codes everywhere; the N x N score tensor never exists in HBM."""
        if B in self._bufs:
            return self._bufs[B]
        C, N, dev = self.C, self.N, self.device
        M = B * N
        Hd = C * self.cfg.mlp_ratio
        K0 = self.cfg.in_chans * self.cfg.patch_size ** 2
        i8 = dict(dtype=torch.int8, device=dev)
        bufs = {
            "patches": torch.empty((B * self.cfg.num_patches, K0), **i8),
            # residual stream (block input / output) and its value after Block.qact2: int16 codes when
            # one of block_input_bw / norm2_in_bw / att_block_out_bw exceeds 8
            "x": torch.empty((M, C), dtype=self.stream_dtype, device=dev),
            "x_mid": torch.empty((M, C), dtype=self.stream_dtype, device=dev),
            "ln": torch.empty((M, C), **i8),         # LayerNorm + QuantAct output (GEMM A operand)
            "qkv": torch.empty((M, 3 * C), **i8),
            "attn": torch.empty((M, C), **i8),
            "h1": torch.empty((M, Hd), **i8),        # fc1 + qact_gelu
            "h2": torch.empty((M, Hd), **i8),        # GELU + mlp.qact1
            "cls": torch.empty((B, C), **i8),
            "logits": torch.empty((B, self.cfg.num_classes), dtype=torch.float32, device=dev),
        }
        self._bufs[B] = bufs
        return bufs

    # ------------------------------------------------------------------ launch sequence
    def _gemm(self, A, lin: _Linear, epilogue, out, name="other", **kw):
        r = ops.gemm(A, lin.w, epilogue, bias=lin.bias, mult=lin.mult, s_in=lin.b_scale, qmin=lin.qmin,
                     qmax=lin.qmax, mult2=lin.mult2, mult_id=lin.mult_id, qmin2=lin.qmin2, qmax2=lin.qmax2,
                     out=out, force_slow=lin.force_slow, z_bound_ok=lin.z_bound_ok, **kw)
        M, K = A.shape
        Nn = lin.w.shape[0]
        idn = kw.get("identity")
        by = (M * K + Nn * K + M * Nn * out.element_size() + Nn * 16 +
              (M * Nn * idn.element_size() if epilogue == EPI_RESIDUAL and idn is not None else 0))
        self._mark("gemm_" + name, 2.0 * M * Nn * K, by)
        return r

    def _ln(self, x, s_in, nm: _Norm, out):
        r = ops.layernorm_requant(nm.kind, x, float(s_in), nm.bias_int, nm.sf, nm.mult, nm.qmin, nm.qmax,
                                  shift=nm.shift, use_int_sqrt=nm.use_int_sqrt, out=out)
        self._mark("layernorm_requant", 0, (1 + x.element_size()) * out.numel() + 16 * self.C)
        return r

    def _enqueue(self, images: torch.Tensor, taps: dict | None = None, blocks: int | None = None) -> torch.Tensor:
        """Enqueue one forward on torch's current stream.  images: float32 [B,3,H,W] on the device.
        taps (tests): name -> int8 code tensor (cloned) under the reference's module names."""
        cfg = self.cfg
        u8 = images.dtype == torch.uint8 and getattr(self, "input_lut", None) is not None
        if self.host_only or not images.is_cuda or (images.dtype != torch.float32 and not u8):
            raise _cabi.IvitKernelError("images must be a float32 CUDA tensor, or uint8 after set_uint8_input "
                                        "(there is no CPU fallback)")
        B = images.shape[0]
        N, C, H, hd = self.N, self.C, self.H, self.hd
        bufs = self._buffers(B)
        self.launches_per_forward = 0

        def tap(name, t, shape=None):
            if taps is not None:
                taps[name] = (t.reshape(shape) if shape else t).clone()

        # stem
        if u8:
            ops.quantize_patches_u8(images, self.input_lut, cfg.patch_size, out=bufs["patches"])
            self._mark("quantize_patches_u8", 0, 2 * images.numel())
        else:
            ops.quantize_patches(images, self.inv_s_input, cfg.patch_size, out=bufs["patches"])
            self._mark("quantize_patches", 0, 5 * images.numel())
        x, x_mid = bufs["x"], bufs["x_mid"]
        self._gemm(bufs["patches"], self.patch, EPI_RESIDUAL, x, name="patch", identity=self.pos_codes, group=cfg.num_patches,
                   out_group_stride=N, out_row_off=1, id_broadcast=True)
        ops.fill_cls_rows(x.view(B, N, C), self.cls_row)
        self._mark("fill_cls_rows", 0, B * C * x.element_size())
        tap("qact1", x, (B, N, C))
        s_x = self.blocks[0].s_in if self.blocks else f32(0)
        nblk = cfg.depth if blocks is None else blocks
        for i, b in enumerate(self.blocks[:nblk]):
            p = f"blocks.{i}."
            self._ln(x, b.s_in, b.norm1, bufs["ln"])
            tap(p + "qact1", bufs["ln"], (B, N, C))
            self._gemm(bufs["ln"], b.qkv, EPI_REQUANT, bufs["qkv"], name="qkv")
            tap(p + "attn.qact1", bufs["qkv"], (B, N, 3 * C))
            ops.attention(bufs["qkv"], B, N, H, hd, b.attn, out=bufs["attn"])
            self._mark("attention_i8", 4.0 * B * H * N * N * hd, 4 * B * N * C)
            tap(p + "attn.qact2", bufs["attn"], (B, N, C))
            self._gemm(bufs["attn"], b.proj, EPI_RESIDUAL, x_mid, name="proj", identity=x)
            tap(p + "qact2", x_mid, (B, N, C))
            self._ln(x_mid, b.s_mid, b.norm2, bufs["ln"])
            tap(p + "qact3", bufs["ln"], (B, N, C))
            # the GELU table pass runs IN PLACE unless the boundary is tapped: most of fc1's 77 MB output is still in L2
            # when the pass reads it, and fc2 finds the codes on the same lines (0.445 -> 0.365 ms per DeiT-S step).
            # Splitting the MLP into L2-sized token ranges was measured and is slower (4.17 / 4.31 / 4.59 ms per step
            # for 2 / 3 / 4 ranges against 4.04).
            h2 = bufs["h2"] if taps is not None else bufs["h1"]
            if self.fuse_gelu:
                # mlp.fc1 -> qact_gelu -> act -> qact1 in ONE launch (layers_quant.py:145-154): the activation-
                # stationary GEMM applies the table on each finished token tile (off by default: measured slower)
                self._gemm(bufs["ln"], b.fc1, EPI_REQUANT, bufs["h1"], name="fc1_gelu", gelu_lut=b.gelu_lut, gelu_out=h2)
                tap(p + "mlp.qact_gelu", bufs["h1"], (B, N, -1))
            else:
                self._gemm(bufs["ln"], b.fc1, EPI_REQUANT, bufs["h1"], name="fc1")
                tap(p + "mlp.qact_gelu", bufs["h1"], (B, N, -1))
                ops.gelu_lut_apply(bufs["h1"], b.gelu_lut, out=h2)
                self._mark("gelu_lut_apply", 0, 2 * bufs["h1"].numel())
            tap(p + "mlp.qact1", h2, (B, N, -1))
            self._gemm(h2, b.fc2, EPI_RESIDUAL, x, name="fc2", identity=x_mid)
            tap(p + "qact4", x, (B, N, C))
            s_x = b.s_out
        # final norm on the class-token rows only (x[:, 0], vit_quant.py:302-304), then the head
        self._ln(x.view(B, N * C)[:, :C], s_x if nblk else self._act_scale("qact1")[0], self.norm, bufs["cls"])
        tap("qact2", bufs["cls"])
        self._gemm(bufs["cls"], self.head, EPI_LOGITS, bufs["logits"], name="head", scale_out=self.head.b_scale)
        return bufs["logits"]

    def forward(self, images: torch.Tensor, taps: dict | None = None, blocks: int | None = None) -> torch.Tensor:
        """Eager launch sequence.  Returns float32 logits [B, classes] (a view of the plan's buffer)."""
        with _device_guard(self.device):
            return self._enqueue(images, taps, blocks)

    __call__ = forward


class GraphedForward:
    """A captured forward: copy images into ``static_in`` (or pass them to ``__call__``), replay."""

    def __init__(self, graph, static_in, static_out, launches):
        self.graph, self.static_in, self.static_out, self.launches = graph, static_in, static_out, launches

    def replay(self) -> torch.Tensor:
        self.graph.replay()
        return self.static_out

    def __call__(self, images: torch.Tensor) -> torch.Tensor:
        self.static_in.copy_(images, non_blocking=True)
        return self.replay()
