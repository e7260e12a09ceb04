"""Seeded synthetic weights / inputs for the ViT-family configs BASELINE.json names.

There is no network for checkpoints or ImageNet, so every test, fixture and
bench run uses these.  The generator is deliberately independent of the
reference's own init order: it emits a plain ``state_dict`` (reference key
names, SURVEY.md section 5 "checkpoint") from a CPU ``torch.Generator``; the
same dict is loaded into the reference model (in the build container, to make
golden fixtures) and into ours (on the GPU box), so both sides see bit-identical
weights.  De-degeneration follows SURVEY.md section 8(d): LN weight ~ U(0.5, 1.5),
biases ~ N(0, 0.1), qkv weight x6 (peaked attention so Shiftmax is exercised).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import torch


@dataclass(frozen=True)
class VitConfig:
    name: str
    embed_dim: int
    depth: int
    num_heads: int
    mlp_ratio: int = 4
    img_size: int = 224
    patch_size: int = 16
    in_chans: int = 3
    num_classes: int = 1000
    # the 8 per-location bit-widths of vit_quant.VisionTransformer.__init__ (vit_quant.py:180-187)
    bitwidths: tuple = (8, 8, 8, 8, 8, 8, 8, 8)
    gelu_type: str = "ivit"
    softmax_type: str = "ivit"
    layernorm_type: str = "ivit"

    @property
    def num_patches(self) -> int:
        return (self.img_size // self.patch_size) ** 2

    @property
    def tokens(self) -> int:
        return self.num_patches + 1

    def with_layers(self, gelu=None, softmax=None, layernorm=None, bitwidths=None) -> "VitConfig":
        from dataclasses import replace
        return replace(self,
                       gelu_type=gelu or self.gelu_type,
                       softmax_type=softmax or self.softmax_type,
                       layernorm_type=layernorm or self.layernorm_type,
                       bitwidths=tuple(bitwidths) if bitwidths is not None else self.bitwidths)


CONFIGS = {
    "deit_tiny": VitConfig("deit_tiny", 192, 12, 3),
    "deit_small": VitConfig("deit_small", 384, 12, 6),
    "deit_base": VitConfig("deit_base", 768, 12, 12),
    "vit_base": VitConfig("vit_base", 768, 12, 12),
    "vit_large": VitConfig("vit_large", 1024, 24, 16),
    # a 2-block, 2-head toy used by fast unit tests (not a reference factory)
    "toy": VitConfig("toy", 128, 2, 2),
}


def synth_vit_state(cfg: VitConfig, seed: int = 0, negate_ln_frac: float = 0.0) -> dict:
    """Deterministic fp32 weights, keyed like the reference ``state_dict`` (weights/biases only)."""
    g = torch.Generator(device="cpu")
    g.manual_seed(1000 + seed)
    C, Hd = cfg.embed_dim, cfg.embed_dim * cfg.mlp_ratio
    P = cfg.patch_size

    def normal(*shape, std):
        return (torch.randn(*shape, generator=g) * std).clamp_(-2.5 * std, 2.5 * std)

    def ln(prefix, sd):
        w = torch.rand(C, generator=g) + 0.5
        if negate_ln_frac > 0:
            flip = torch.rand(C, generator=g) < negate_ln_frac
            w = torch.where(flip, -w, w)
        sd[prefix + ".weight"] = w
        sd[prefix + ".bias"] = normal(C, std=0.1)

    sd: dict = {}
    sd["cls_token"] = normal(1, 1, C, std=0.02)
    sd["pos_embed"] = normal(1, cfg.tokens, C, std=0.02)
    sd["patch_embed.proj.weight"] = normal(C, cfg.in_chans, P, P, std=0.02)
    sd["patch_embed.proj.bias"] = normal(C, std=0.1)
    for i in range(cfg.depth):
        b = f"blocks.{i}."
        ln(b + "norm1", sd)
        sd[b + "attn.qkv.weight"] = normal(3 * C, C, std=0.02) * 6.0
        sd[b + "attn.qkv.bias"] = normal(3 * C, std=0.1)
        sd[b + "attn.proj.weight"] = normal(C, C, std=0.02)
        sd[b + "attn.proj.bias"] = normal(C, std=0.1)
        ln(b + "norm2", sd)
        sd[b + "mlp.fc1.weight"] = normal(Hd, C, std=0.02)
        sd[b + "mlp.fc1.bias"] = normal(Hd, std=0.1)
        sd[b + "mlp.fc2.weight"] = normal(C, Hd, std=0.02)
        sd[b + "mlp.fc2.bias"] = normal(C, std=0.1)
    ln("norm", sd)
    sd["head.weight"] = normal(cfg.num_classes, C, std=0.02)
    sd["head.bias"] = normal(cfg.num_classes, std=0.1)
    return sd


def synth_images(batch: int, cfg: VitConfig, seed: int = 0) -> torch.Tensor:
    """ImageNet-normalised images are ~N(0,1); fp32 [B,3,H,W] on CPU."""
    g = torch.Generator(device="cpu")
    g.manual_seed(2000 + seed)
    return torch.randn(batch, cfg.in_chans, cfg.img_size, cfg.img_size, generator=g)


# ---------------------------------------------------------------------------------------------
# Calibrated activation ranges.  The frozen plan is a function of (weights, x_min/x_max buffers);
# for the synthetic weights above those buffers were produced by the reference's own calibration
# pass (scripts/inference.py:33-91, one forward of `synth_images(4, cfg, seed=1)` with
# running_stat=True, then freeze_model) and are committed in tests/golden/vit_<case>.npz by
# oracle/make_golden.py.  A real deployment reads the same buffers from a reference checkpoint.
# name -> (config, gelu, softmax, layernorm, negate_ln_frac[, bitwidths of the 8 activation locations])
CASES = {
    "toy_ivit": ("toy", "ivit", "ivit", "ivit", 0.0),
    "toy_ibert": ("toy", "ibert", "ibert", "ibert", 0.0),
    "toy_mixed": ("toy", "ivit", "ibert", "ivit", 0.05),
    "toy_ibert_intsqrt": ("toy", "ibert", "ibert", "ibert_use-int-sqrt_true", 0.0),
    "deit_tiny_ivit": ("deit_tiny", "ivit", "ivit", "ivit", 0.0),
    "deit_tiny_ibert": ("deit_tiny", "ibert", "ibert", "ibert", 0.0),
    "deit_small_ivit": ("deit_small", "ivit", "ivit", "ivit", 0.0),
    "vit_base_mixed": ("vit_base", "ivit", "ibert", "ivit", 0.0),
    "deit_base_ibert": ("deit_base", "ibert", "ibert", "ibert", 0.0),
    "vit_large_ivit": ("vit_large", "ivit", "ivit", "ivit", 0.0),
    # 16-bit activation locations (vit_quant.py:180-187)
    "toy_ivit_w16": ("toy", "ivit", "ivit", "ivit", 0.0, (16,) * 8),
    "toy_ibert_w16": ("toy", "ibert", "ibert", "ibert", 0.0, (16,) * 8),
    "toy_mixed_bw": ("toy", "ivit", "ibert", "ivit", 0.05, (16, 8, 8, 16, 16, 8, 16, 8)),
    "toy_ivit_sm16": ("toy", "ivit", "ivit", "ivit", 0.0, (8, 8, 8, 8, 16, 8, 8, 8)),
    "deit_tiny_ibert_w16": ("deit_tiny", "ibert", "ibert", "ibert", 0.0, (16,) * 8),
    "deit_tiny_ivit_w16": ("deit_tiny", "ivit", "ivit", "ivit", 0.0, (16,) * 8),
    "deit_base_ibert_w16": ("deit_base", "ibert", "ibert", "ibert", 0.0, (16,) * 8),    # BASELINE.json configs[4]
    # the `ppoly` registry entries (layer_selection.py:116-131)
    "toy_ppoly": ("toy", "ppoly", "ppoly", "ivit", 0.0),
    "toy_ppoly_args": ("toy", "ppoly_scale-bits_20_seg_8_backend_float", "ppoly_exp-bits_12_seg_8", "ibert", 0.0),
    "deit_tiny_ppoly": ("deit_tiny", "ppoly", "ppoly", "ivit", 0.0),
}


def golden_path(case: str) -> str:
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    return os.path.join(root, "tests", "golden", f"vit_{case}.npz")


def load_case(case: str):
    """Returns (cfg, state) for a committed case: seeded weights + the calibrated range buffers
    (``*.x_min``, ``*.x_max``, ``*.shift``) as float32 numpy arrays, keyed like the reference
    state_dict.  Also returns the raw fixture (reference logits, tap hashes) as third element."""
    import numpy as np
    cfg_name, g, s, l, neg = CASES[case][:5]
    bw = CASES[case][5] if len(CASES[case]) > 5 else None
    cfg = CONFIGS[cfg_name].with_layers(gelu=g, softmax=s, layernorm=l, bitwidths=bw)
    state = {k: v.numpy() for k, v in synth_vit_state(cfg, seed=0, negate_ln_frac=neg).items()}
    fix = np.load(golden_path(case))
    for k in fix.files:
        if k.startswith("range/"):
            state[k[len("range/"):]] = fix[k].astype(np.float32)
        elif k.startswith("fixed/"):                    # ppoly layers: stored integer coefficients (state_dict buffers)
            state[k[len("fixed/"):]] = fix[k]
    return cfg, state, fix


# ---------------------------------------------------------------------------------------------
# Swin (swin_quant.py): BASELINE.json configs[3].  The reference hard-codes the ivit layer types
# (IntLayerNorm / IntSoftmax / IntGELU aliases) and the bit-widths (16-bit residual stream).
@dataclass(frozen=True)
class SwinConfig:
    name: str
    embed_dim: int = 96
    depths: tuple = (2, 2, 6, 2)
    num_heads: tuple = (3, 6, 12, 24)
    window_size: int = 7
    mlp_ratio: int = 4
    img_size: int = 224
    patch_size: int = 4
    in_chans: int = 3
    num_classes: int = 1000

    @property
    def grid(self) -> int:
        return self.img_size // self.patch_size

    def stage_dim(self, i: int) -> int:
        return self.embed_dim * 2 ** i

    def stage_res(self, i: int) -> int:
        return self.grid // 2 ** i

    def block_window_shift(self, i: int, j: int):
        """swin_quant.py:200-203: no partition / shift when the resolution is not larger than the window."""
        res, ws = self.stage_res(i), self.window_size
        shift = 0 if j % 2 == 0 else ws // 2
        if res <= ws:
            return res, 0
        return ws, shift


SWIN_CONFIGS = {
    "swin_tiny": SwinConfig("swin_tiny"),
    # 2-stage toy: 56x56 image -> 14x14 tokens (4 windows, shifted block with mask) -> 7x7 (one window)
    "swin_toy": SwinConfig("swin_toy", embed_dim=96, depths=(2, 2), num_heads=(3, 6), img_size=56, num_classes=100),
}

SWIN_CONFIGS.update({
    # Swin-B widths (C = 128 -> 256, LayerNorm over 512 channels in PatchMerging, 4 / 8 heads of 32)
    "swin_toy_b": SwinConfig("swin_toy_b", embed_dim=128, depths=(2, 2), num_heads=(4, 8), img_size=56, num_classes=100),
    # 4 x 4 windows (16 tokens, shift 2) on an 8 x 8 token grid, then one 4 x 4 window
    "swin_toy_w4": SwinConfig("swin_toy_w4", embed_dim=96, depths=(2, 2), num_heads=(3, 6), window_size=4, img_size=32,
                              num_classes=100),
})

SWIN_CASES = {"swin_toy": "swin_toy", "swin_tiny": "swin_tiny", "swin_toy_b": "swin_toy_b", "swin_toy_w4": "swin_toy_w4"}


def synth_swin_state(cfg: SwinConfig, seed: int = 0) -> dict:
    """Deterministic fp32 weights keyed like swin_quant.SwinTransformer.state_dict() (weights only)."""
    g = torch.Generator(device="cpu")
    g.manual_seed(3000 + seed)

    def normal(*shape, std):
        return (torch.randn(*shape, generator=g) * std).clamp_(-2.5 * std, 2.5 * std)

    def ln(prefix, C, sd):
        sd[prefix + ".weight"] = torch.rand(C, generator=g) + 0.5
        sd[prefix + ".bias"] = normal(C, std=0.1)

    sd: dict = {}
    C0, P = cfg.embed_dim, cfg.patch_size
    sd["patch_embed.proj.weight"] = normal(C0, cfg.in_chans, P, P, std=0.1)
    sd["patch_embed.proj.bias"] = normal(C0, std=0.1)
    ln("patch_embed.norm", C0, sd)
    ws = cfg.window_size
    for i, depth in enumerate(cfg.depths):
        C, nH = cfg.stage_dim(i), cfg.num_heads[i]
        for j in range(depth):
            b = f"layers.{i}.blocks.{j}."
            w, _ = cfg.block_window_shift(i, j)
            ln(b + "norm1", C, sd)
            sd[b + "attn.qkv.weight"] = normal(3 * C, C, std=0.02) * 6.0
            sd[b + "attn.qkv.bias"] = normal(3 * C, std=0.1)
            sd[b + "attn.relative_position_bias_table"] = normal((2 * w - 1) ** 2, nH, std=0.5)
            sd[b + "attn.proj.weight"] = normal(C, C, std=0.02)
            sd[b + "attn.proj.bias"] = normal(C, std=0.1)
            ln(b + "norm2", C, sd)
            sd[b + "mlp.fc1.weight"] = normal(C * cfg.mlp_ratio, C, std=0.02)
            sd[b + "mlp.fc1.bias"] = normal(C * cfg.mlp_ratio, std=0.1)
            sd[b + "mlp.fc2.weight"] = normal(C, C * cfg.mlp_ratio, std=0.02)
            sd[b + "mlp.fc2.bias"] = normal(C, std=0.1)
        if i < len(cfg.depths) - 1:
            d = f"layers.{i}.downsample."
            ln(d + "norm", 4 * C, sd)
            sd[d + "reduction.weight"] = normal(2 * C, 4 * C, std=0.02)
    Cl = cfg.stage_dim(len(cfg.depths) - 1)
    ln("norm", Cl, sd)
    sd["head.weight"] = normal(cfg.num_classes, Cl, std=0.02)
    sd["head.bias"] = normal(cfg.num_classes, std=0.1)
    return sd


def load_swin_case(case: str):
    """(cfg, state incl. calibrated ranges, fixture) of a committed Swin case (tests/golden/swin_<case>.npz)."""
    import os
    import numpy as np
    cfg = SWIN_CONFIGS[SWIN_CASES[case]]
    state = {k: v.numpy() for k, v in synth_swin_state(cfg, seed=0).items()}
    fix = np.load(os.path.join(os.path.dirname(golden_path("x")), f"swin_{case}.npz"))
    for k in fix.files:
        if k.startswith("range/"):
            state[k[len("range/"):]] = fix[k].astype(np.float32)
        elif k.startswith("fixed/"):                    # ppoly layers: stored integer coefficients (state_dict buffers)
            state[k[len("fixed/"):]] = fix[k]
    return cfg, state, fix
