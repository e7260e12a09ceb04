"""Drop-in for the reference's `models/quantization_utils` (its export list, __init__.py:1-4), plus
the three `Int*` aliases that `swin_quant.py:11` imports but the reference never exported."""
import torch
import torch.nn.functional as F

from .carrier import QTensor
from .floatmods import FloatGELU, FloatLayerNorm, FloatSoftmax
from .iostats import (attach_io_stat_hooks, clear_io_stats, disable_io_stats, enable_io_stats, get_io_stats_df,
                      save_io_stats_df)
from .modules import (IBERTIntGELU, IBERTIntLayerNorm, IBERTIntSoftmax, IVITIntGELU, IVITIntLayerNorm,
                      IVITIntSoftmax, QuantAct, QuantConv2d, QuantLinear, QuantMatMul)
from .ppolymods import PPolyIntGELU, PPolyIntSoftmax
from .registry import get_gelu, get_layernorm, get_softmax

IntLayerNorm, IntSoftmax, IntGELU = IVITIntLayerNorm, IVITIntSoftmax, IVITIntGELU


def softmax(x, dim: int, onnx_trace: bool = False):
    """quant_utils.py:263-267."""
    if isinstance(x, QTensor):
        x = x.dequantize()
    return F.softmax(x.float(), dim=dim) if onnx_trace else F.softmax(x, dim=dim, dtype=torch.float32)


__all__ = ["QuantLinear", "QuantAct", "QuantConv2d", "QuantMatMul", "attach_io_stat_hooks", "save_io_stats_df",
           "enable_io_stats", "disable_io_stats", "clear_io_stats", "softmax", "IVITIntGELU", "IVITIntSoftmax",
           "IVITIntLayerNorm", "IBERTIntGELU", "IBERTIntSoftmax", "IBERTIntLayerNorm", "get_gelu", "get_softmax",
           "get_layernorm", "IntLayerNorm", "IntSoftmax", "IntGELU", "QTensor", "FloatGELU", "FloatSoftmax",
           "FloatLayerNorm", "PPolyIntGELU", "PPolyIntSoftmax"]
