"""Checkpoint -> calibrated -> frozen lifecycle of a model built on the quantization_utils module API.

Mirrors the helpers of the reference's ``scripts/inference.py`` (``load_model`` :94-224,
``calibrate_model`` :33-91) and ``models/model_utils.py`` (``freeze_model`` / ``unfreeze_model``) for
ANY caller that is made of this package's modules -- the reference's own ``vit_quant.py`` imported
over ``ivit_b200.quantization_utils`` or a caller of one's own.  Data
loading (ImageFolder, transforms) stays with the caller: ``calibrate_model`` takes an iterable of
image batches.

The running-stat calibration forward is device-agnostic torch code (it is not the hot path); after
``freeze_model`` the modules exchange integer codes and launch the sm_100a kernels, CUDA only.
"""
from __future__ import annotations

from typing import Iterable, Mapping

import numpy as np
import torch
from torch import nn

from .quantization_utils import modules as _m

_FREEZABLE = (_m.QuantAct, _m.QuantLinear, _m.QuantConv2d, _m.QuantMatMul, _m.IVITIntLayerNorm, _m.IBERTIntLayerNorm,
              _m.IVITIntGELU, _m.IBERTIntGELU, _m.IVITIntSoftmax, _m.IBERTIntSoftmax)


class _FusedPlanForward:
    """``model.forward`` of a frozen Swin model: the integer forward as the fused static plan.

    The frozen module path carries swin_quant.py's glue too (torch.roll, window partition, the bias gather,
    the float mask add: ``carrier.py``), but it launches per module and moves the window permutations with copy
    kernels; ``FrozenSwinPlan`` folds them into index maps of the window-attention kernel.  Same kernels,
    bit-identical logits (tests/test_swin_gpu.py), built once per device from the model's own ``state_dict``."""

    def __init__(self, model: nn.Module):
        self._model = model
        self._plans = {}
        # the plans hold constants derived from the state_dict: a later load_state_dict starts over
        self._hook = model.register_load_state_dict_post_hook(lambda module, incompatible: self._plans.clear())

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        from . import _cabi
        from .api import swin_config_from_model
        from .swin_plan import FrozenSwinPlan
        if not x.is_cuda:
            raise _cabi.IvitKernelError("a frozen model runs the integer kernels: CUDA tensors only (no CPU fallback); "
                                        "use unfreeze_model() for the float-simulated calibration mode")
        key = (x.device.type, x.device.index)
        if key not in self._plans:
            self._plans[key] = FrozenSwinPlan(swin_config_from_model(self._model), self._model.state_dict(), x.device)
        return self._plans[key].forward(x.float().contiguous()).clone()


def _is_swin(model: nn.Module) -> bool:
    return hasattr(model, "layers") and not hasattr(model, "blocks") and hasattr(model, "patch_embed") and \
        all(hasattr(st, "blocks") for st in model.layers)


def freeze_model(model: nn.Module, fused_plan: bool = True) -> nn.Module:
    """model_utils.py ``freeze_model``: every module that has ``fix()`` stops updating its running
    statistics; from here on the forward is the integer path.  A Swin model additionally gets its
    forward routed to the fused plan (see ``_FusedPlanForward``, the fast path: ~90 launches instead of
    one or more per module); ``fused_plan=False`` -- or the reference's own ``freeze_model``, which only
    calls ``fix()`` -- leaves ``swin_quant.py``'s forward in place and it runs module by module on the
    integer carriers, bit-identical.  ``unfreeze_model`` undoes either."""
    model.eval()
    for mod in model.modules():                 # model_utils.py:13-15: every sub-module that provides fix()
        if hasattr(mod, "fix") and callable(mod.fix):
            mod.fix()
    if _is_swin(model) and fused_plan:
        model.__dict__["forward"] = _FusedPlanForward(model)
    return model


def unfreeze_model(model: nn.Module) -> nn.Module:
    """model_utils.py ``unfreeze_model``: back to running-stat (calibration / QAT) mode."""
    for mod in model.modules():                 # (ppoly modules keep their fitted coefficients: model_utils.py:23)
        if isinstance(mod, _FREEZABLE) and hasattr(mod, "unfix"):
            mod.unfix()
    if isinstance(model.__dict__.get("forward"), _FusedPlanForward):
        model.__dict__["forward"]._hook.remove()
        del model.__dict__["forward"]
    return model


def load_checkpoint(model: nn.Module, checkpoint: str | Mapping, strict: bool = False, map_location="cpu"):
    """inference.py:186-205: accept a path or a dict, the new (``{'model': state_dict, ...}``) and the
    old (bare ``state_dict``) formats, and 0-d buffers where the model holds shape-[1] ones.
    Returns ``(missing_keys, unexpected_keys)``."""
    ck = torch.load(checkpoint, map_location=map_location) if isinstance(checkpoint, (str, bytes)) or hasattr(
        checkpoint, "__fspath__") else checkpoint
    weights = dict(ck["model"]) if "model" in ck and isinstance(ck["model"], Mapping) else dict(ck)
    own = model.state_dict()
    for name, value in list(weights.items()):
        value = torch.as_tensor(value)
        if name in own and value.shape == torch.Size([]) and own[name].shape == torch.Size([1]):
            value = value.unsqueeze(0)
        weights[name] = value
    res = model.load_state_dict(weights, strict=strict)
    return list(res.missing_keys), list(res.unexpected_keys)


def calibrate_model(model: nn.Module, batches: Iterable[torch.Tensor], device=None) -> int:
    """inference.py:33-91 without the data loading: eval mode, running statistics on, one forward
    per batch under ``no_grad``.  Returns the number of batches seen."""
    model.eval()
    unfreeze_model(model)
    n = 0
    with torch.no_grad():
        for x in batches:
            if isinstance(x, (tuple, list)):      # (images, labels) pairs of a DataLoader
                x = x[0]
            model(x.to(device) if device is not None else x)
            n += 1
    return n


def load_model(model: nn.Module, checkpoint: str | Mapping, calibration_batches: Iterable[torch.Tensor] | None = None,
               device=None, strict_load: bool = False) -> nn.Module:
    """inference.py ``load_model`` for an already constructed model: weights in, optional
    calibration (skipped with ``strict_load``, which then requires every range buffer to come from
    the checkpoint), frozen model out."""
    load_checkpoint(model, checkpoint, strict=strict_load)
    if device is not None:
        model.to(device)
    model.eval()
    if not strict_load and calibration_batches is not None:
        calibrate_model(model, calibration_batches, device)
    return freeze_model(model)


def tvm_params(model: nn.Module) -> dict:
    """The arrays ``TVM_benchmark/convert_model.py:12-66`` (``save_params``) pulls out of the
    state_dict -- every ``*.weight_integer`` as int8, every ``*.bias_integer`` as int32 -- keyed by
    their checkpoint names, plus every ``*scaling_factor`` (``load_qconfig`` :69-148) as float32.
    They exist after one calibration (or frozen) forward, with the reference's values."""
    out = {}
    for key, t in model.state_dict().items():
        a = t.detach().cpu().numpy()
        if key.endswith("weight_integer"):
            out[key] = a.astype(np.int8)
        elif key.endswith("bias_integer"):
            out[key] = a.astype(np.int32)
        elif key.endswith("scaling_factor"):
            out[key] = a.astype(np.float32)
    return out
