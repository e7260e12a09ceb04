"""--gelu / --softmax / --layernorm selection (layer_selection.py:116-236): the `ivit` and `ibert`
entries and the `base_arg_val_arg_val` name grammar that injects constructor keywords
(`ibert_use-int-sqrt_true` -> IBERTIntLayerNorm(use_int_sqrt=True)), the `float` diagnostics
(true float op + quantised output, torch ops in every mode) and the `ppoly` piecewise-polynomial
GELU / softmax (`ppoly_scale-bits_20_seg_8` ...)."""
from __future__ import annotations

from .floatmods import FloatGELU, FloatLayerNorm, FloatSoftmax
from .ppolymods import PPolyIntGELU, PPolyIntSoftmax
from .modules import (IBERTIntGELU, IBERTIntLayerNorm, IBERTIntSoftmax, IVITIntGELU, IVITIntLayerNorm,
                      IVITIntSoftmax)

GELU_REGISTRY = {"float": FloatGELU, "ivit": IVITIntGELU, "ibert": IBERTIntGELU, "ppoly": PPolyIntGELU}
SOFTMAX_REGISTRY = {"float": FloatSoftmax, "ivit": IVITIntSoftmax, "ibert": IBERTIntSoftmax, "ppoly": PPolyIntSoftmax}
LN_REGISTRY = {"float": FloatLayerNorm, "ivit": IVITIntLayerNorm, "ibert": IBERTIntLayerNorm}
_NOT_BUILT = ()


def _coerce(text: str):
    low = text.lower()
    if low in ("true", "false"):
        return low == "true"
    if text.isdigit():
        return int(text)
    try:
        return float(text)
    except ValueError:
        return text


def parse_layer_name(name: str):
    """'base_k1_v1_k2_v2' -> ('base', {'k1': v1, 'k2': v2}); hyphens in keys become underscores."""
    parts = name.lower().split("_")
    if len(parts) < 3:
        return name.lower(), {}
    params = {}
    for key, val in zip(parts[1:-1:2], parts[2::2]):
        params[key.replace("-", "_")] = _coerce(val)
    return parts[0], params


def _select(registry: dict, name: str, what: str):
    base, params = parse_layer_name(name)
    if base not in registry:
        if base in _NOT_BUILT or name.lower() in _NOT_BUILT:
            raise NotImplementedError(f"{what} type '{name}': only the float|ivit|ibert entries are built on this path")
        return registry[name.lower()]          # KeyError for unknown names, like the reference
    cls = registry[base]
    if not params:
        return cls

    class Parameterized(cls):                  # constructor keywords injected as defaults
        def __init__(self, *args, **kwargs):
            for k, v in params.items():
                kwargs.setdefault(k, v)
            super().__init__(*args, **kwargs)

    Parameterized.__name__ = f"Parameterized{cls.__name__}"
    return Parameterized


def get_gelu(name: str):
    return _select(GELU_REGISTRY, name, "GELU")


def get_softmax(name: str):
    return _select(SOFTMAX_REGISTRY, name, "softmax")


def get_layernorm(name: str):
    return _select(LN_REGISTRY, name, "layernorm")
