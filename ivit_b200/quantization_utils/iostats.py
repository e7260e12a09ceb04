"""Per-layer I/O statistics collector (quant_modules.py:17-125): forward hooks that log integer
extrema and shapes of what every module returns.  Works on both carriers; integer carriers that
are still pending (not yet launched) are logged by shape only, so the hooks never break a fusion."""
from __future__ import annotations

from functools import partial

import torch
import torch.nn as nn

from .carrier import QTensor

_LAYER_IO_STATS: list = []
_ENABLED = True


def _codes_minmax(t, scale):
    if isinstance(t, QTensor):
        if t.codes is None:
            return None, None
        c = t.materialize()
        return float(c.min()), float(c.max())
    if scale is None or not torch.is_tensor(t):
        return None, None
    s = scale
    if s.numel() > 1 and s.numel() == t.shape[-1]:
        s = s.view(*([1] * (t.dim() - 1)), -1)
    q = t.detach() / s
    return float(q.min()), float(q.max())


def _collect(module, inputs, output, module_name):
    if not _ENABLED:
        return
    try:
        rec = {"layer": module_name, "type": type(module).__name__}
        if isinstance(output, tuple) and len(output) == 2:
            y, s = output
            rec["out_shape"] = tuple(y.shape)
            rec["min_out_int"], rec["max_out_int"] = _codes_minmax(y, s)
        if inputs:
            x = inputs[0]
            rec["in_shape"] = tuple(x.shape) if hasattr(x, "shape") else None
            s_in = inputs[1] if len(inputs) > 1 and torch.is_tensor(inputs[1]) else None
            rec["min_in_int"], rec["max_in_int"] = _codes_minmax(x, s_in)
        _LAYER_IO_STATS.append(rec)
    except Exception:
        pass                                   # statistics must never break a forward (reference: :79-81)


def attach_io_stat_hooks(model: nn.Module):
    for name, module in model.named_modules():
        if module is model:
            continue
        module.register_forward_hook(partial(_collect, module_name=name))


def enable_io_stats():
    global _ENABLED
    _ENABLED = True


def disable_io_stats():
    global _ENABLED
    _ENABLED = False


def clear_io_stats():
    _LAYER_IO_STATS.clear()


def get_io_stats_df():
    import pandas as pd
    return pd.DataFrame(_LAYER_IO_STATS)


def save_io_stats_df(path: str = "io_stats.pkl", to_csv: bool = False):
    df = get_io_stats_df()
    df.to_pickle(path)
    if to_csv:
        df.to_csv(path.rsplit(".", 1)[0] + ".csv", index=False)
    return df
