"""TEST INFRASTRUCTURE ONLY (build container only: needs /root/reference).

Builds the *reference's own* VisionTransformer (``vit_quant.py``) with our seeded
synthetic weights, runs its calibrate -> freeze lifecycle
(``scripts/inference.py:33-91`` calibration pass with ``running_stat=True``,
``model_utils.py:5`` freeze_model) and dumps what its modules produce, through
plain forward hooks.  Used by ``oracle/make_golden.py`` and by the CPU tests that
pin ``oracle/ivit_oracle.py`` / ``oracle/vit_oracle.py`` against the reference.
"""
from __future__ import annotations

import numpy as np
import torch

from .ref_shim import load_reference

_FACTORY = {
    "deit_tiny": "deit_tiny_patch16_224",
    "deit_small": "deit_small_patch16_224",
    "deit_base": "deit_base_patch16_224",
    "vit_base": "vit_base_patch16_224",
    "vit_large": "vit_large_patch16_224",
}


def build_reference_vit(cfg, state: dict):
    """Instantiate the reference model for `cfg` and load `state` (weights, optionally buffers)."""
    ref = load_reference()
    bw = cfg.bitwidths
    kw = dict(patch_embed_bw=bw[0], pos_encoding_bw=bw[1], block_input_bw=bw[2], attention_out_bw=bw[3],
              softmax_bw=bw[4], mlp_out_bw=bw[5], norm2_in_bw=bw[6], att_block_out_bw=bw[7],
              gelu_type=cfg.gelu_type, softmax_type=cfg.softmax_type, layernorm_type=cfg.layernorm_type)
    if cfg.name in _FACTORY:
        model = getattr(ref, _FACTORY[cfg.name])(pretrained=False, **kw)
    else:  # e.g. the "toy" config: same class, custom sizes
        import sys
        vq = sys.modules[ref.__name__ + ".vit_quant"]      # VisionTransformer is not in models.__all__
        model = vq.VisionTransformer(patch_size=cfg.patch_size, embed_dim=cfg.embed_dim, depth=cfg.depth,
                                      num_heads=cfg.num_heads, mlp_ratio=cfg.mlp_ratio, qkv_bias=True,
                                      img_size=cfg.img_size, num_classes=cfg.num_classes, **kw)
    missing, unexpected = model.load_state_dict({k: torch.as_tensor(v) for k, v in state.items()}, strict=False)
    assert not unexpected, unexpected
    bad = [k for k in missing if k.endswith((".weight", ".bias")) or k in ("cls_token", "pos_embed")]
    assert not bad, bad
    model.eval()
    return model


def calibrate_and_freeze(model, calib_images: torch.Tensor):
    """One calibration forward with running_stat=True, then freeze."""
    ref = load_reference()
    with torch.no_grad():
        model(calib_images)
    ref.freeze_model(model)
    return model


def act_ranges(model) -> dict:
    """The calibrated buffers that define every frozen scale: *.x_min / *.x_max (+ ibert LN *.shift)."""
    out = {}
    for k, v in model.state_dict().items():
        if k.endswith((".x_min", ".x_max", ".shift")):
            out[k] = v.detach().clone().reshape(-1).numpy().astype(np.float32)
    return out


def run_with_taps(model, images: torch.Tensor, names=None):
    """Forward with hooks; returns (logits ndarray, {module name: (carrier ndarray, scale ndarray)})."""
    taps = {}
    handles = []

    def mk(name):
        def hook(mod, inp, out):
            if isinstance(out, tuple) and len(out) == 2 and torch.is_tensor(out[1]):
                taps[name] = (out[0].detach().numpy().copy(), out[1].detach().numpy().copy())
        return hook

    containers = {"Block", "Attention", "Mlp", "PatchEmbed", "VisionTransformer", "ModuleList"}
    for name, mod in model.named_modules():
        if not name or type(mod).__name__ in containers:
            continue
        if names is not None and name not in names:
            continue
        handles.append(mod.register_forward_hook(mk(name)))
    with torch.no_grad():
        y = model(images)
    for h in handles:
        h.remove()
    return y.numpy(), taps


def build_reference_swin(cfg, state: dict):
    """The reference's own SwinTransformer (swin_quant.py:419) with our seeded weights."""
    from functools import partial

    from .ref_shim import give_biasless_linears_a_zero_bias, load_reference_swin

    sw = load_reference_swin()
    ref = load_reference()
    import sys
    qu = sys.modules[ref.__name__ + ".quantization_utils"]
    model = sw.SwinTransformer(img_size=cfg.img_size, patch_size=cfg.patch_size, in_chans=cfg.in_chans,
                               num_classes=cfg.num_classes, embed_dim=cfg.embed_dim, depths=cfg.depths,
                               num_heads=cfg.num_heads, window_size=cfg.window_size, mlp_ratio=float(cfg.mlp_ratio),
                               norm_layer=partial(qu.IntLayerNorm, eps=1e-6))
    give_biasless_linears_a_zero_bias(model)
    missing, unexpected = model.load_state_dict({k: torch.as_tensor(v) for k, v in state.items()}, strict=False)
    assert not unexpected, unexpected
    bad = [k for k in missing if k.endswith((".weight", "relative_position_bias_table")) or
           (k.endswith(".bias") and "reduction" not in k)]
    assert not bad, bad
    model.eval()
    return model


def run_swin_with_taps(model, images: torch.Tensor):
    """Forward with hooks on every leaf quantisation module; returns (logits, {name: (carrier, scale)})."""
    taps, handles = {}, []

    def mk(name):
        def hook(mod, inp, out):
            if isinstance(out, tuple) and len(out) == 2 and torch.is_tensor(out[1]):
                taps[name] = (out[0].detach().numpy().copy(), out[1].detach().numpy().copy())
        return hook

    leaf = {"QuantAct", "QuantLinear", "QuantConv2d", "QuantMatMul", "IVITIntLayerNorm", "IVITIntGELU", "IVITIntSoftmax"}
    for name, mod in model.named_modules():
        if type(mod).__name__ in leaf:
            handles.append(mod.register_forward_hook(mk(name)))
    with torch.no_grad():
        y = model(images)
    for h in handles:
        h.remove()
    return y.numpy(), taps
