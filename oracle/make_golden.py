"""TEST INFRASTRUCTURE ONLY (build container only: needs /root/reference).

Generates the committed fixtures under ``tests/golden/`` by RUNNING THE UNMODIFIED REFERENCE
(``oracle/ref_shim.py`` shims, no source edits) on seeded synthetic weights and inputs:

  * ``vit_<case>.npz``   one per model case: the calibrated ``x_min`` / ``x_max`` / ``shift``
    buffers of every QuantAct (what ``scripts/inference.py:33-91`` calibration + ``freeze_model``
    leave behind), the reference logits, SHA-256 of the integer codes and of the scale at EVERY
    module boundary (forward hooks, reference module names), and the full codes of a few taps.
  * ``modules.npz``      per-function known answers: fixedpoint_mul / QuantAct, IntLayerNorm,
    IntGELU (the reference's own 256-code sweep, scripts/gelu_approx_analysis.py:73-77),
    IntSoftmax, for both ivit and ibert, on seeded inputs.

The fixtures travel to the GPU box, where /root/reference does not exist; there they (a) pin
``oracle/ivit_oracle.py`` / ``oracle/vit_oracle.py`` (tests/test_oracle_golden.py) and (b) provide
the calibrated activation ranges the frozen CUDA plan is built from.

Usage (build container):  python -m oracle.make_golden [--only NAME]
"""
from __future__ import annotations

import argparse
import hashlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from ivit_b200.synthetic import CONFIGS, synth_images, synth_vit_state  # noqa: E402
from oracle import ref_harness as RH  # noqa: E402
from oracle.ref_shim import load_reference  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
f32 = np.float32

# name -> (config, gelu, softmax, layernorm, bitwidths, batch, negate_ln_frac, taps stored in full)
CASES = {
    "toy_ivit": ("toy", "ivit", "ivit", "ivit", None, 2, 0.0, ("blocks.0.attn.qact2", "blocks.1.qact4")),
    "toy_ibert": ("toy", "ibert", "ibert", "ibert", None, 2, 0.0, ("blocks.0.attn.qact2", "blocks.1.qact4")),
    "toy_mixed": ("toy", "ivit", "ibert", "ivit", None, 2, 0.05, ("blocks.1.qact4",)),
    "toy_ibert_intsqrt": ("toy", "ibert", "ibert", "ibert_use-int-sqrt_true", None, 2, 0.0, ()),
    # the `ppoly` registry entries (layer_selection.py:116-131): piecewise-polynomial GELU and softmax
    "toy_ppoly": ("toy", "ppoly", "ppoly", "ivit", None, 2, 0.0, ("blocks.0.attn.qact2", "blocks.1.qact4")),
    "toy_ppoly_args": ("toy", "ppoly_scale-bits_20_seg_8_backend_float", "ppoly_exp-bits_12_seg_8", "ibert", None, 2, 0.0, ()),
    "deit_tiny_ppoly": ("deit_tiny", "ppoly", "ppoly", "ivit", None, 2, 0.0, ("qact2",)),
    # BASELINE.json configs[0]: DeiT-Tiny INT8 ivit, batch 8
    "deit_tiny_ivit": ("deit_tiny", "ivit", "ivit", "ivit", None, 8, 0.0, ("qact2",)),
    "deit_tiny_ibert": ("deit_tiny", "ibert", "ibert", "ibert", None, 2, 0.0, ("qact2",)),
    # configs[1]: DeiT-Small ivit (reference run on the first 4 images of the seeded batch)
    "deit_small_ivit": ("deit_small", "ivit", "ivit", "ivit", None, 4, 0.0, ("qact2",)),
    # configs[2]: ViT-Base, mixed layer selection
    "vit_base_mixed": ("vit_base", "ivit", "ibert", "ivit", None, 2, 0.0, ("qact2",)),
    "deit_base_ibert": ("deit_base", "ibert", "ibert", "ibert", None, 2, 0.0, ("qact2",)),
    # the largest factory of vit_quant.py (:390-406): ViT-Large, 24 blocks of width 1024
    "vit_large_ivit": ("vit_large", "ivit", "ivit", "ivit", None, 2, 0.0, ("qact2",)),
    # 16-bit activation locations (vit_quant.py:180-187; scripts/inference.py:297 --bitwidth 16): all eight
    # at 16 bit, and a mix (patch_embed, pos, block_input, attention_out, softmax, mlp_out, norm2_in, block_out)
    "toy_ivit_w16": ("toy", "ivit", "ivit", "ivit", (16,) * 8, 2, 0.0, ("blocks.0.attn.qact2", "blocks.1.qact4")),
    "toy_ibert_w16": ("toy", "ibert", "ibert", "ibert", (16,) * 8, 2, 0.0, ("blocks.0.attn.qact2", "blocks.1.qact4")),
    "toy_mixed_bw": ("toy", "ivit", "ibert", "ivit", (16, 8, 8, 16, 16, 8, 16, 8), 2, 0.05, ("blocks.1.qact4",)),
    "toy_ivit_sm16": ("toy", "ivit", "ivit", "ivit", (8, 8, 8, 8, 16, 8, 8, 8), 2, 0.0, ("blocks.0.attn.qact2",)),
    # BASELINE.json configs[4] family: 16-bit ibert layer types on a DeiT (tiny here: CPU-sized oracle run)
    "deit_tiny_ibert_w16": ("deit_tiny", "ibert", "ibert", "ibert", (16,) * 8, 2, 0.0, ("qact2",)),
    "deit_tiny_ivit_w16": ("deit_tiny", "ivit", "ivit", "ivit", (16,) * 8, 2, 0.0, ("qact2",)),
    # configs[4]: DeiT-Base, ibert layer types, --bitwidth 16
    "deit_base_ibert_w16": ("deit_base", "ibert", "ibert", "ibert", (16,) * 8, 2, 0.0, ("qact2",)),
}
CALIB_BATCH = 4


def case_config(name):
    cfg_name, g, s, l, bw, batch, neg, full = CASES[name]
    cfg = CONFIGS[cfg_name].with_layers(gelu=g, softmax=s, layernorm=l, bitwidths=bw)
    return cfg, batch, neg, full


def sha(a) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def tap_codes(y, s) -> np.ndarray:
    """Integer codes a downstream QuantAct would recover: round(y / s) (quant_utils.py:220)."""
    return np.rint(np.asarray(y, dtype=f32) / np.asarray(s, dtype=f32)).astype(np.int64)


def make_vit_case(name: str) -> dict:
    cfg, batch, neg, full = case_config(name)
    sd = synth_vit_state(cfg, seed=0, negate_ln_frac=neg)
    model = RH.build_reference_vit(cfg, sd)
    RH.calibrate_and_freeze(model, synth_images(CALIB_BATCH, cfg, seed=1))
    ranges = RH.act_ranges(model)
    x = synth_images(batch, cfg, seed=0)
    logits, taps = RH.run_with_taps(model, x)
    out = {"logits": logits.astype(f32), "meta": np.array([name, cfg.name, cfg.gelu_type, cfg.softmax_type,
                                                           cfg.layernorm_type, str(batch), str(neg)])}
    for k, v in ranges.items():
        out["range/" + k] = v
    # ppoly layers: the integer coefficients the frozen modules fitted on their first batch (ppoly_modules.py:57-108)
    for k, v in model.state_dict().items():
        if ".fixed_" in k:
            out["fixed/" + k] = v.detach().cpu().numpy()
    names, code_sha, scale_sha = [], [], []
    for k, (y, s) in taps.items():
        names.append(k)
        code_sha.append(sha(tap_codes(y, s)))
        scale_sha.append(sha(np.asarray(s, dtype=f32)))
        if k in full:
            c = tap_codes(y, s)
            out["codes/" + k] = c.astype(np.int8 if np.abs(c).max() < 128 else np.int32)
            out["scale/" + k] = np.asarray(s, dtype=f32)
    out["tap_names"] = np.array(names)
    out["tap_code_sha"] = np.array(code_sha)
    out["tap_scale_sha"] = np.array(scale_sha)
    return out


# Swin (BASELINE.json configs[3]): name -> (config, batch, calibration batch, taps stored in full)
SWIN_CASES = {
    "swin_toy": ("swin_toy", 2, 4, ("layers.0.blocks.1.attn.qact3", "layers.1.blocks.1.qact4")),
    "swin_tiny": ("swin_tiny", 2, 2, ("qact3",)),
    "swin_toy_b": ("swin_toy_b", 2, 4, ("layers.1.blocks.1.qact4",)),
    "swin_toy_w4": ("swin_toy_w4", 2, 4, ("layers.0.blocks.1.attn.qact3",)),
}


def make_swin_case(name: str) -> dict:
    """The reference's own swin_quant.SwinTransformer (shimmed import, oracle/ref_shim.load_reference_swin)
    on the seeded weights: calibrate, freeze, forward with a hook on every quantisation module."""
    from ivit_b200.synthetic import SWIN_CONFIGS, synth_swin_state
    cfg_name, batch, calib, full = SWIN_CASES[name]
    cfg = SWIN_CONFIGS[cfg_name]
    model = RH.build_reference_swin(cfg, synth_swin_state(cfg, seed=0))
    RH.calibrate_and_freeze(model, synth_images(calib, cfg, seed=1))
    ranges = RH.act_ranges(model)
    logits, taps = RH.run_swin_with_taps(model, synth_images(batch, cfg, seed=0))
    out = {"logits": logits.astype(f32), "meta": np.array([name, cfg.name, str(batch)])}
    for k, v in ranges.items():
        out["range/" + k] = v
    # memory layout of every LayerNorm input in the reference run (decides torch's float32 summation order)
    names, code_sha, scale_sha = [], [], []
    for k, (y, s) in taps.items():
        names.append(k)
        code_sha.append(sha(tap_codes(y, s)))
        scale_sha.append(sha(np.asarray(s, dtype=f32)))
        if k in full:
            c = tap_codes(y, s)
            out["codes/" + k] = c.astype(np.int8 if np.abs(c).max() < 128 else np.int32)
            out["scale/" + k] = np.asarray(s, dtype=f32)
    out["tap_names"], out["tap_code_sha"], out["tap_scale_sha"] = np.array(names), np.array(code_sha), np.array(scale_sha)
    return out


def make_modules() -> dict:
    """Known answers of the reference's module functions on seeded inputs."""
    ref = load_reference()
    qu = sys.modules[ref.__name__ + ".quantization_utils"]
    g = torch.Generator().manual_seed(7)
    out = {}

    def carrier(shape, scale, lo=-128, hi=127, spread=40.0):
        q = torch.clamp(torch.round(torch.randn(*shape, generator=g) * spread), lo, hi)
        return (q * scale).float(), q

    with torch.no_grad():
        # --- QuantAct / fixedpoint_mul: scalar and per-channel predecessor scale, identity branch
        s_in = torch.tensor([0.0173])
        x, q = carrier((3, 50, 96), s_in)
        idn, qi = carrier((3, 50, 96), torch.tensor([0.031]))
        for bits in (8, 16):
            act = qu.QuantAct(bits)
            act.x_min.fill_(-1.9)
            act.x_max.fill_(2.3)
            act.fix()
            y, s = act(x, s_in)
            out[f"qact{bits}/x"], out[f"qact{bits}/s_in"] = x.numpy(), s_in.numpy()
            out[f"qact{bits}/codes"] = tap_codes(y.numpy(), s.numpy()).astype(np.int32)
            y, s = act(x, s_in, idn, torch.tensor([0.031]))
            out[f"qact{bits}/id"] = idn.numpy()
            out[f"qact{bits}/codes_id"] = tap_codes(y.numpy(), s.numpy()).astype(np.int32)
        s_pc = (torch.rand(96, generator=g) * 2e-3 + 1e-4)
        s_pc[::7] *= -1        # negative per-channel scales (LayerNorm weights < 0, hazard H5)
        zi = torch.round(torch.randn(3, 50, 96, generator=g) * 3e4)
        xpc = (zi * s_pc).float()
        act = qu.QuantAct(8)
        act.x_min.fill_(-40.0)
        act.x_max.fill_(55.0)
        act.fix()
        y, s = act(xpc, s_pc)
        out["qact_pc/x"], out["qact_pc/s_in"] = xpc.numpy(), s_pc.numpy()
        out["qact_pc/codes"] = tap_codes(y.numpy(), s.numpy()).astype(np.int32)
        # no predecessor scale: plain symmetric quantiser
        xf = torch.randn(2, 3, 32, 32, generator=g) * 1.2
        act = qu.QuantAct(8)
        act.x_min.fill_(float(xf.min()))
        act.x_max.fill_(float(xf.max()))
        act.fix()
        y, s = act(xf)
        out["qact_in/x"] = xf.numpy()
        out["qact_in/codes"] = tap_codes(y.numpy(), s.numpy()).astype(np.int32)

        # --- LayerNorm (both kinds), C = 192
        C = 192
        s_ln = torch.tensor([0.0371])
        xl, _ = carrier((2, 67, C), s_ln, spread=35.0)
        w = torch.rand(C, generator=g) + 0.5
        w[::11] *= -1
        b = torch.randn(C, generator=g) * 0.1
        out["ln/x"], out["ln/s_in"], out["ln/w"], out["ln/b"] = xl.numpy(), s_ln.numpy(), w.numpy(), b.numpy()
        for kind, cls, kw in (("ivit", qu.IVITIntLayerNorm, {}), ("ibert", qu.IBERTIntLayerNorm, {}),
                              ("ibert_intsqrt", qu.IBERTIntLayerNorm, {"use_int_sqrt": True})):
            ln = cls(C, **kw)
            ln.weight.copy_(w)
            ln.bias.copy_(b)
            ln.eval()
            if hasattr(ln, "fix"):
                ln.fix()
            y, s = ln(xl, s_ln)
            out[f"ln/{kind}/y_int"] = (y / s).numpy().astype(f32)
            out[f"ln/{kind}/scale"] = s.numpy().astype(f32)

        # --- GELU: the reference's own 256-code sweep (scripts/gelu_approx_analysis.py:73-77)
        for sc in (0.08, 0.0371, 0.0123):
            s_g = torch.tensor([sc])
            xg = (torch.arange(-128, 128).float() * s_g).reshape(1, 1, 256)
            for kind, cls in (("ivit", qu.IVITIntGELU), ("ibert", qu.IBERTIntGELU)):
                y, s = cls()(xg, s_g)
                out[f"gelu/{kind}/{sc}/y_int"] = (y / s).numpy().astype(f32)
                out[f"gelu/{kind}/{sc}/scale"] = s.numpy().astype(f32).reshape(-1)
        # random rows (row max matters for ShiftGELU)
        s_g = torch.tensor([0.0371])
        xg, _ = carrier((2, 9, 512), s_g, hi=100, spread=30.0)
        out["gelu/rows/x"] = xg.numpy()
        y, s = qu.IVITIntGELU()(xg, s_g)
        out["gelu/rows/ivit_y_int"] = (y / s).numpy().astype(f32)

        # --- softmax
        s_sm = torch.tensor([0.0625])
        xs, _ = carrier((2, 2, 197, 197), s_sm, spread=30.0)
        out["softmax/x"], out["softmax/s_in"] = xs.numpy(), s_sm.numpy()
        y, s = qu.IVITIntSoftmax(8)(xs, s_sm)
        out["softmax/ivit/codes"] = tap_codes(y.numpy(), s.numpy()).astype(np.uint8)
        sm = qu.IBERTIntSoftmax(8)
        sm.eval()
        y, s = sm(xs, s_sm)       # first call calibrates the inner QuantAct(16)
        sm.act.fix()
        y, s = sm(xs, s_sm)
        out["softmax/ibert/act_x_min"] = sm.act.x_min.numpy().reshape(-1)
        out["softmax/ibert/act_x_max"] = sm.act.x_max.numpy().reshape(-1)
        out["softmax/ibert/codes"] = tap_codes(y.numpy(), s.numpy()).astype(np.uint8)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default=None)
    args = ap.parse_args()
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    torch.set_num_threads(os.cpu_count() or 1)
    if args.only in (None, "modules"):
        np.savez_compressed(os.path.join(GOLDEN_DIR, "modules.npz"), **make_modules())
        print("modules.npz")
    for name in CASES:
        if args.only not in (None, name):
            continue
        d = make_vit_case(name)
        path = os.path.join(GOLDEN_DIR, f"vit_{name}.npz")
        np.savez_compressed(path, **d)
        print(path, os.path.getsize(path) // 1024, "KiB", len(d["tap_names"]), "taps")
    for name in SWIN_CASES:
        if args.only not in (None, name):
            continue
        d = make_swin_case(name)
        path = os.path.join(GOLDEN_DIR, f"swin_{name}.npz")
        np.savez_compressed(path, **d)
        print(path, os.path.getsize(path) // 1024, "KiB", len(d["tap_names"]), "taps")


if __name__ == "__main__":
    main()
