"""Frozen static plan of the Swin integer forward (swin_quant.py; BASELINE.json configs[3]).

Same idea as ``plan.FrozenVitPlan``: after ``freeze_model`` every scale, dyadic multiplier, integer
weight / bias, LayerNorm ``bias_int``, GELU table and the quantised relative-position bias are functions
of (weights, ``x_min`` / ``x_max`` buffers) only; they are computed once on the host (``hostmath``) and
``SwinTransformer.forward`` (swin_quant.py:536-563) is replayed as launches of the kernels behind
``include/ivit_b200.h``.

Kernel fusion map (reference modules -> one launch):
  qact_input + im2col (4x4 patches)                         -> ivit_quantize_patches_f32_i8
  patch_embed.proj + qact_before_norm                       -> ivit_gemm_i8 (REQUANT)
  patch_embed.norm + patch_embed.qact                       -> ivit_layernorm_requant_strided (int8 rows)
  qact1 (16 bit)                                            -> ivit_requant_codes
  norm1 + Block.qact1 | norm2 + qact3 | norm + qact2        -> ivit_layernorm_requant_i16_i8 / _strided (stage 0)
  attn.qkv + attn.qact1                                     -> ivit_gemm_i8 (REQUANT)
  roll + window_partition + matmul_1 * scale + qact_attn1 + qact2(+bias) + mask + softmax + matmul_2 + qact3
      + window_reverse + roll back                          -> ivit_window_attention_i8
  attn.proj + attn.qact4(16) + Block.qact2(16, identity)    -> ivit_gemm_i8 (RESIDUAL, int16 stream)
  mlp.fc1 + qact_gelu | mlp.act + mlp.qact1                 -> ivit_gemm_i8 (REQUANT) | ivit_gelu_lut_apply_i8
  mlp.fc2 + mlp.qact2 + Block.qact4(16, identity)           -> ivit_gemm_i8 (RESIDUAL, int16 stream)
  PatchMerging slices + cat | norm + qact1 | reduction + qact2 -> ivit_patch_merge_gather | LN | GEMM (REQUANT, int16 out)
  avgpool + qact3 | head                                    -> ivit_avgpool_requant_i8 | ivit_gemm_i8 (LOGITS)

The residual stream is int16 (swin_quant.py:110,214,222,475 hard-code QuantAct(16)).  Stage 0's
LayerNorms use torch's outer-reduction summation order (DESIGN.md hazard H11).  No CPU path.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _cabi, hostmath as hm, ops
from ._cabi import EPI_LOGITS, EPI_REQUANT, EPI_RESIDUAL, AttentionParams, WindowAttentionParams
from .plan import _device_guard, GraphCaptureMixin, _Linear, _Norm, _np
from .synthetic import SwinConfig

f32 = np.float32


def relative_position_index(ws: int) -> np.ndarray:
    """swin_quant.py:80-95."""
    coords = np.stack(np.meshgrid(np.arange(ws), np.arange(ws), indexing="ij")).reshape(2, -1)
    rel = (coords[:, :, None] - coords[:, None, :]).transpose(1, 2, 0).copy()
    rel[:, :, 0] += ws - 1
    rel[:, :, 1] += ws - 1
    rel[:, :, 0] *= 2 * ws - 1
    return rel.sum(-1)


def shift_window_mask(res: int, ws: int, shift: int) -> np.ndarray:
    """float32 [nW, ws*ws, ws*ws] of 0 / -100: SwinTransformerBlock.attn_mask (swin_quant.py:223-247) -- region labels
    of the rolled image (three bands per axis), partitioned into windows; pairs from different regions get -100."""
    band = np.where(np.arange(res) < res - ws, 0, np.where(np.arange(res) < res - shift, 1, 2))
    lab = 3 * band[:, None] + band[None, :]
    n = res // ws
    mw = lab.reshape(n, ws, n, ws).transpose(0, 2, 1, 3).reshape(n * n, ws * ws)
    return np.where(mw[:, None, :] != mw[:, :, None], f32(-100.0), f32(0.0)).astype(f32)


def mask_always_clamps(s2, qmin2: int, qmax2: int) -> bool:
    """float32 restatement of the kernel's masked score: x = fl(fl(c s) - 100) / s against the smallest
    possible row maximum fl(fl(qmin s) / s); int_exp_shift (ivit_modules.py:150-162) clamps at 15 x0."""
    s = f32(s2)
    x0 = np.floor(f32(-1.0) / s)
    xm = f32(f32(f32(f32(qmax2) * s) + f32(-100.0)) / s)
    xmin = f32(f32(f32(qmin2) * s) / s)
    d = f32(xm - xmin)                                     # largest possible (masked score - row maximum)
    t = f32(f32(d + np.floor(d * f32(0.5))) - np.floor(d * f32(0.0625)))      # :151, non-decreasing in d
    return bool(np.isfinite(x0) and d < 0 and t <= f32(15.0) * x0 - f32(2.0))


class FrozenSwinPlan(GraphCaptureMixin):
    def __init__(self, cfg: SwinConfig, state: dict, device="cuda", host_only: bool = False):
        self.cfg, self.host_only = cfg, host_only
        self.device = torch.device("cpu" if host_only else device)
        if not host_only and self.device.type != "cuda":
            raise _cabi.IvitKernelError("FrozenSwinPlan needs a CUDA device (there is no CPU fallback)")
        if not host_only:
            _cabi.load()
        self.sd = {k: _np(v) for k, v in state.items() if not k.endswith(("weight_integer", "bias_integer"))}
        self._bufs, self._keep, self._graphs = {}, [], {}
        self.launches_per_forward = 0
        self.trace = None
        with _device_guard(self.device):        # constants, tables and launches on the plan's device, whatever is current
            self._build()

    # ------------------------------------------------------------------ host constants
    def _act_scale(self, name: str, bits: int = 8) -> np.ndarray:
        return hm.sym_scale(bits, self.sd[name + ".x_min"], self.sd[name + ".x_max"]).reshape(1)

    def _dev(self, a, dtype):
        return ops.dev(a, dtype, self.device)

    def _linear(self, name, s_x, s_out, bits=8) -> _Linear:
        p = hm.linear_params(self.sd[name + ".weight"], self.sd.get(name + ".bias"), s_x)
        mult, slow = None, False
        if s_out is not None:
            m = hm.dyadic_mult(p["b_scale"], s_out)
            slow = bool(np.abs(m).max() >= 256.0) or bool(np.abs(p["b_int_f32"] if p["b_int_f32"] is not None
                                                               else 0).max() >= 2 ** 30)
            mult = self._dev(m.reshape(-1), torch.float64)
        lo, hi = hm.qrange(bits)
        zb = np.abs(p["w_int_f32"]).sum(axis=1).max() * 128.0 + np.abs(p["b_int"].astype(np.float64)).max()
        return _Linear(z_bound_ok=bool(zb < 2 ** 22) and not slow, w=self._dev(p["w_int"], torch.int8),
                       bias=self._dev(p["b_int"], torch.int32), b_scale=self._dev(p["b_scale"], torch.float32), mult=mult,
                       qmin=lo, qmax=hi, force_slow=slow, s_out=(s_out[0] if s_out is not None else f32(0)))

    def _residual(self, lin: _Linear, s_mid, s_id, s_out, bits):
        lin.mult2 = float(hm.dyadic_mult(s_mid, s_out).reshape(-1)[0])
        lin.mult_id = float(hm.dyadic_mult(s_id, s_out).reshape(-1)[0])
        lin.qmin2, lin.qmax2 = hm.qrange(bits)
        lin.s_out2 = s_out[0]
        if max(abs(lin.mult2), abs(lin.mult_id)) >= 2 ** 20:
            lin.force_slow = True
        return lin

    def _norm(self, name, C, s_out) -> _Norm:
        p = hm.layernorm_params(self.sd[name + ".weight"], self.sd[name + ".bias"], C)
        mult = hm.dyadic_mult(p["sf"], s_out).reshape(-1)
        return _Norm(kind=0, bias_int=self._dev(p["bias_int"], torch.float32), sf=self._dev(p["sf"], torch.float32),
                     mult=self._dev(mult, torch.float64), s_out=s_out[0])

    def _gelu_lut(self, s_in, s_out):
        s = f32(s_in[0])
        s_g = np.array([s * f32(1 / 2 ** 7)], dtype=f32)                   # ivit_modules.py:121-124
        mult = float(hm.dyadic_mult(s_g, s_out).reshape(-1)[0])
        if self.host_only:
            return None
        lut, s_dev = ops.gelu_lut_build(0, float(s), 8, mult, -128, 127, self.device)
        if f32(s_dev) != s_g[0]:
            raise _cabi.IvitKernelError("GELU output scale: host and library disagree")
        return lut

    def _window_attention(self, p, C, nH, ws, s_qkv):
        """swin_quant.py:131-160: scales of matmul_1, * scale, qact_attn1, qact_table, qact2, softmax, matmul_2, qact3."""
        a = WindowAttentionParams()
        hd = C // nH
        s_mm1 = (s_qkv * s_qkv).astype(f32)
        s_scaled = (s_mm1 * f32(hd ** -0.5)).astype(f32)                   # :135-136
        s_a1 = self._act_scale(p + "qact_attn1")
        a.mult_attn = float(hm.dyadic_mult(s_scaled, s_a1).reshape(-1)[0])
        a.qmin_attn, a.qmax_attn = -128, 127
        s_tab = self._act_scale(p + "qact_table")
        tab = hm.quantize_tensor(self.sd[p + "relative_position_bias_table"], 8, s_tab)        # :140-141
        N = ws * ws
        bias = tab[relative_position_index(ws).reshape(-1)].reshape(N, N, nH).transpose(2, 0, 1)   # :142-145
        bias_dev = self._dev(np.ascontiguousarray(bias), torch.int8)
        self._keep.append(bias_dev)
        a.bias_codes = bias_dev.data_ptr() if not self.host_only else None
        s2 = self._act_scale(p + "qact2")
        a.mult2 = float(hm.dyadic_mult(s_a1, s2).reshape(-1)[0])
        a.mult_bias = float(hm.dyadic_mult(s_tab, s2).reshape(-1)[0])
        a.qmin2, a.qmax2 = -128, 127
        # identity branch of qact2, batch independent: rint(bias * m 2^-e) in float64 (exact product: 8 + 32 bits)
        terms = np.rint(np.ascontiguousarray(bias).astype(np.float64) * a.mult_bias)
        if np.abs(terms).max() < 2 ** 31 and not self.host_only:
            terms_dev = self._dev(terms.astype(np.int64), torch.int32)
            self._keep.append(terms_dev)
            a.bias_terms = terms_dev.data_ptr()
        a.s2 = float(s2[0])
        a.output_bit = 8
        # exp term tabulated over (row-max code, code) by the library's own arithmetic; a masked score must
        # provably hit int_exp_shift's clamp for the table path to be taken with a shift mask
        a.mask_clamps = int(mask_always_clamps(s2[0], a.qmin2, a.qmax2))
        if not self.host_only:
            tp = AttentionParams()
            tp.kind, tp.output_bit, tp.s_attn = 0, 8, float(s2[0])
            self._keep.append(ops.softmax_table_build(tp, self.device, collapse=False))
            a.exp_table = tp.exp_table
            a.sat_k, a.sat_e = tp.sat_k, tp.sat_e
        s_sm = np.array([1 / 2 ** 7], dtype=f32)                           # ivit_modules.py:177
        s_mm2 = (s_sm * s_qkv).astype(f32)
        s3 = self._act_scale(p + "qact3")
        a.mult_out = float(hm.dyadic_mult(s_mm2, s3).reshape(-1)[0])
        a.qmin_out, a.qmax_out = -128, 127
        return a, s3

    def _build(self):
        cfg = self.cfg
        C0 = cfg.embed_dim
        self.s_input = self._act_scale("qact_input")
        self.inv_s_input = float(f32(1.0) / self.s_input[0])
        s_bn = self._act_scale("patch_embed.qact_before_norm")
        self.patch = self._linear("patch_embed.proj", self.s_input, s_bn)
        s_pe = self._act_scale("patch_embed.qact")
        self.patch_norm = self._norm("patch_embed.norm", C0, s_pe)
        self.s_bn = s_bn[0]
        s_x = self._act_scale("qact1", 16)
        self.stem_mult = self._dev(hm.dyadic_mult(s_pe, s_x).reshape(-1), torch.float64)
        self.stem_s_in = self._dev(s_pe, torch.float32)
        self.stages = []
        for i, depth in enumerate(cfg.depths):
            C, nH = cfg.stage_dim(i), cfg.num_heads[i]
            blocks = []
            for j in range(depth):
                p = f"layers.{i}.blocks.{j}."
                ws, shift = cfg.block_window_shift(i, j)
                b = {"ws": ws, "shift": shift, "s_in": s_x[0]}
                s_n1 = self._act_scale(p + "qact1")
                b["norm1"] = self._norm(p + "norm1", C, s_n1)
                s_qkv = self._act_scale(p + "attn.qact1")
                b["qkv"] = self._linear(p + "attn.qkv", s_n1, s_qkv)
                b["attn"], s3 = self._window_attention(p + "attn.", C, nH, ws, s_qkv)
                s_q4 = self._act_scale(p + "attn.qact4", 16)
                s_mid = self._act_scale(p + "qact2", 16)
                b["proj"] = self._residual(self._linear(p + "attn.proj", s3, s_q4, 16), s_q4, s_x, s_mid, 16)
                s_n2 = self._act_scale(p + "qact3")
                b["norm2"] = self._norm(p + "norm2", C, s_n2)
                s_g = self._act_scale(p + "mlp.qact_gelu")
                b["fc1"] = self._linear(p + "mlp.fc1", s_n2, s_g)
                s_m1 = self._act_scale(p + "mlp.qact1")
                b["gelu_lut"] = self._gelu_lut(s_g, s_m1)
                s_m2 = self._act_scale(p + "mlp.qact2")
                s_o = self._act_scale(p + "qact4", 16)
                b["fc2"] = self._residual(self._linear(p + "mlp.fc2", s_m1, s_m2), s_m2, s_mid, s_o, 16)
                b["s_mid"] = s_mid[0]
                blocks.append(b)
                s_x = s_o
            st = {"blocks": blocks, "C": C, "nH": nH, "res": cfg.stage_res(i)}
            if i < len(cfg.depths) - 1:
                d = f"layers.{i}.downsample."
                s_d1 = self._act_scale(d + "qact1")
                st["merge_norm"] = self._norm(d + "norm", 4 * C, s_d1)
                st["merge_s_in"] = s_x[0]
                s_d2 = self._act_scale(d + "qact2")
                st["reduction"] = self._linear(d + "reduction", s_d1, s_d2)      # bias=False -> zero bias
                s_x = s_d2
            self.stages.append(st)
        self.s_final_in = s_x[0]
        Cl = cfg.stage_dim(len(cfg.depths) - 1)
        s_f = self._act_scale("qact2")
        self.norm = self._norm("norm", Cl, s_f)
        s_p = self._act_scale("qact3")
        self.pool_s_in = float(s_f[0])
        self.pool_mult = float(hm.dyadic_mult(s_f, s_p).reshape(-1)[0])
        self.head = self._linear("head", s_p, None)
        if not self.host_only:
            torch.cuda.synchronize(self.device)

    # ------------------------------------------------------------------ buffers
    def _buffers(self, B: int) -> dict:
        if B in self._bufs:
            return self._bufs[B]
        cfg, dev = self.cfg, self.device
        L0, C0 = cfg.grid ** 2, cfg.embed_dim
        M0 = B * L0
        K0 = cfg.in_chans * cfg.patch_size ** 2
        e = lambda shape, dt: torch.empty(shape, dtype=dt, device=dev)
        # element counts shrink by 2 per stage (L / 4, C * 2): stage 0 sizes bound every later stage
        bufs = {
            "patches": e((M0, K0), torch.int8),
            "pe": e((M0, C0), torch.int8), "pe2": e((M0, C0), torch.int8),
            "x": e((M0 * C0,), torch.int16), "x_mid": e((M0 * C0,), torch.int16), "xm": e((M0 * C0,), torch.int16),
            "ln": e((M0 * C0,), torch.int8), "qkv": e((M0 * 3 * C0,), torch.int8), "attn": e((M0 * C0,), torch.int8),
            "h1": e((M0 * C0 * cfg.mlp_ratio,), torch.int8), "h2": e((M0 * C0 * cfg.mlp_ratio,), torch.int8),
            "pool": e((B, cfg.stage_dim(len(cfg.depths) - 1)), torch.int8),
            "logits": e((B, cfg.num_classes), torch.float32),
        }
        self._bufs[B] = bufs
        return bufs

    # ------------------------------------------------------------------ launch sequence
    def _gemm(self, A, lin: _Linear, epilogue, out, name="other", **kw):
        r = ops.gemm(A, lin.w, epilogue, bias=lin.bias, mult=lin.mult, s_in=lin.b_scale, qmin=lin.qmin, qmax=lin.qmax,
                     mult2=lin.mult2, mult_id=lin.mult_id, qmin2=lin.qmin2, qmax2=lin.qmax2, out=out,
                     force_slow=lin.force_slow, z_bound_ok=lin.z_bound_ok, **kw)
        M, K = A.shape
        Nn = lin.w.shape[0]
        idn = kw.get("identity")
        by = (M * K + Nn * K + M * Nn * out.element_size() + Nn * 16 +
              (M * Nn * idn.element_size() if idn is not None else 0))
        self._mark("gemm_" + name, 2.0 * M * Nn * K, by)
        return r

    def _ln(self, x, s_in, nm: _Norm, out, strided_tokens=0):
        if strided_tokens:
            r = ops.layernorm_requant_strided(0, x, strided_tokens, float(s_in), nm.bias_int, nm.sf, nm.mult, nm.qmin,
                                              nm.qmax, out=out)
        else:
            r = ops.layernorm_requant(0, x, float(s_in), nm.bias_int, nm.sf, nm.mult, nm.qmin, nm.qmax, out=out)
        self._mark("layernorm_requant", 0, (1 + x.element_size()) * out.numel() + 16 * x.shape[-1])
        return r

    def forward(self, images: torch.Tensor, taps: dict | None = None) -> torch.Tensor:
        """Eager launch sequence.  Returns float32 logits [B, classes] (a view of the plan's buffer)."""
        with _device_guard(self.device):
            return self._enqueue(images, taps)

    __call__ = forward

    def _enqueue(self, images: torch.Tensor, taps: dict | None = None) -> torch.Tensor:
        cfg = self.cfg
        if self.host_only or not images.is_cuda or images.dtype != torch.float32:
            raise _cabi.IvitKernelError("images must be a float32 CUDA tensor (there is no CPU fallback)")
        B = images.shape[0]
        bufs = self._buffers(B)
        self.launches_per_forward = 0

        def tap(name, t, shape):
            if taps is not None:
                taps[name] = t.reshape(shape).clone()

        L, C = cfg.grid ** 2, cfg.embed_dim
        M = B * L
        ops.quantize_patches(images, self.inv_s_input, cfg.patch_size, out=bufs["patches"])
        self._mark("quantize_patches", 0, 5 * images.numel())
        self._gemm(bufs["patches"], self.patch, EPI_REQUANT, bufs["pe"], name="patch")
        tap("patch_embed.qact_before_norm", bufs["pe"], (B, L, C))
        self._ln(bufs["pe"], self.s_bn, self.patch_norm, bufs["pe2"], strided_tokens=L)
        tap("patch_embed.qact", bufs["pe2"], (B, L, C))
        x = bufs["x"][:M * C].view(M, C)
        ops.requant(bufs["pe2"], self.stem_s_in, self.stem_mult, -32768, 32767, out=x, out_bits=16)
        self._mark("requant_codes", 0, 3 * x.numel())
        tap("qact1", x, (B, L, C))
        cur, other = "x", "xm"
        for i, st in enumerate(self.stages):
            C, nH, res = st["C"], st["nH"], st["res"]
            L = res * res
            M = B * L
            Hd = C * cfg.mlp_ratio
            x = bufs[cur][:M * C].view(M, C)
            x_mid = bufs["x_mid"][:M * C].view(M, C)
            ln = bufs["ln"][:M * C].view(M, C)
            qkv = bufs["qkv"][:M * 3 * C].view(M, 3 * C)
            attn = bufs["attn"][:M * C].view(M, C)
            h1, h2 = bufs["h1"][:M * Hd].view(M, Hd), bufs["h2"][:M * Hd].view(M, Hd)
            strided = L if i == 0 else 0
            for j, b in enumerate(st["blocks"]):
                p = f"layers.{i}.blocks.{j}."
                self._ln(x, b["s_in"], b["norm1"], ln, strided)
                tap(p + "qact1", ln, (B, L, C))
                self._gemm(ln, b["qkv"], EPI_REQUANT, qkv, name="qkv")
                ops.window_attention(qkv, B, res, b["ws"], b["shift"], nH, C // nH, b["attn"], out=attn)
                self._mark("window_attention_i8", 4.0 * M * b["ws"] ** 2 * C, 4 * M * C)
                tap(p + "attn.qact3.imgorder", attn, (B, L, C))
                self._gemm(attn, b["proj"], EPI_RESIDUAL, x_mid, name="proj", identity=x)
                tap(p + "qact2", x_mid, (B, L, C))
                self._ln(x_mid, b["s_mid"], b["norm2"], ln, strided)
                tap(p + "qact3", ln, (B, L, C))
                self._gemm(ln, b["fc1"], EPI_REQUANT, h1, name="fc1")
                tap(p + "mlp.qact_gelu", h1, (B, L, Hd))
                g_out = h2 if taps is not None else h1          # in place unless tapped: fc1's rows are still in L2
                ops.gelu_lut_apply(h1, b["gelu_lut"], out=g_out)
                self._mark("gelu_lut_apply", 0, 2 * h1.numel())
                tap(p + "mlp.qact1", g_out, (B, L, Hd))
                self._gemm(g_out, b["fc2"], EPI_RESIDUAL, x, name="fc2", identity=x_mid)
                tap(p + "qact4", x, (B, L, C))
            if "reduction" in st:
                d = f"layers.{i}.downsample."
                Mq = M // 4
                xm = bufs[other][:M * C].view(Mq, 4 * C)
                ops.patch_merge_gather(x, B, res, C, out=xm)
                self._mark("patch_merge_gather", 0, 4 * x.numel())
                ln4 = bufs["ln"][:Mq * 4 * C].view(Mq, 4 * C)
                self._ln(xm, st["merge_s_in"], st["merge_norm"], ln4)
                tap(d + "qact1", ln4, (B, L // 4, 4 * C))
                nx = bufs[cur][:Mq * 2 * C].view(Mq, 2 * C)          # x is dead after the gather
                self._gemm(ln4, st["reduction"], EPI_REQUANT, nx, name="merge")    # 8-bit codes into the int16 stream
                tap(d + "qact2", nx, (B, L // 4, 2 * C))
        C = self.stages[-1]["C"]
        L = self.stages[-1]["res"] ** 2
        M = B * L
        x = bufs[cur][:M * C].view(M, C)
        ln = bufs["ln"][:M * C].view(M, C)
        self._ln(x, self.s_final_in, self.norm, ln)
        tap("qact2", ln, (B, L, C))
        ops.avgpool_requant(ln, B, L, C, self.pool_s_in, self.pool_mult, -128, 127, out=bufs["pool"])
        self._mark("avgpool_requant", 0, ln.numel() + B * C)
        tap("qact3", bufs["pool"], (B, C))
        self._gemm(bufs["pool"], self.head, EPI_LOGITS, bufs["logits"], name="head", scale_out=self.head.b_scale)
        return bufs["logits"]
