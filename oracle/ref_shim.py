"""TEST INFRASTRUCTURE ONLY -- never imported by the product path.

Imports the *unmodified* reference (lionnus/I-ViT, mounted read-only at
/root/reference) on a CPU-only box.  The reference cannot be imported as-is
(SURVEY.md section 0): `ppoly_backend.py:8` imports matplotlib at module top level and
six sites call `.cuda(device=x.device)` unconditionally (`quant_utils.py:89`,
`:174-175`; `ivit_modules.py:33`, `:121`, `:176`; `ibert_modules.py:126`,
`:317`).  Two shims, no source edits:

  * a stub `matplotlib` / `matplotlib.pyplot` in `sys.modules`
  * `torch.Tensor.cuda` becomes a no-op while the reference runs on CPU

This module only exists in the build container: `/root/reference` is absent on
the GPU box, so nothing under `tests/ -m gpu`, `smoke()` or `bench.py` imports
it.  It is used by `oracle/make_golden.py` (fixture generator) and by the
`-m "not gpu"` tests that pin the numpy oracle against the real reference.
"""
from __future__ import annotations

import contextlib
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("IVIT_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "models", "vit_quant.py"))


def _install_stubs() -> None:
    if "matplotlib" not in sys.modules:
        mpl = types.ModuleType("matplotlib")
        plt = types.ModuleType("matplotlib.pyplot")
        mpl.pyplot = plt
        sys.modules["matplotlib"] = mpl
        sys.modules["matplotlib.pyplot"] = plt
    if "tkinter" not in sys.modules:           # swin_quant.py:2 `from tkinter import X`
        tk = types.ModuleType("tkinter")
        tk.X = "x"
        sys.modules["tkinter"] = tk


_REF = None


def load_reference():
    """Return the reference's `models` package (ViT/DeiT factories + quantization_utils)."""
    global _REF
    if _REF is not None:
        return _REF
    if not reference_available():
        raise RuntimeError(f"reference not mounted at {REFERENCE_ROOT}")
    import torch

    _install_stubs()
    # CPU run of code that hard-codes `.cuda(device=...)`: only where there is no CUDA device, so that a box with
    # both the reference and a GPU keeps the real method for the `-m gpu` tests of the same process.
    if not torch.cuda.is_available():
        torch.Tensor.cuda = lambda self, device=None, *a, **k: self  # type: ignore[assignment]
    # The reference's top-level package is called `models`; keep it from
    # colliding with anything of ours by importing it under its own sys.path entry.
    sys.path.insert(0, REFERENCE_ROOT)
    try:
        import importlib

        _REF = importlib.import_module("models")
    finally:
        sys.path.remove(REFERENCE_ROOT)
    return _REF


@contextlib.contextmanager
def torch_threads(n: int | None = None):
    import torch

    old = torch.get_num_threads()
    torch.set_num_threads(n or os.cpu_count() or 1)
    try:
        yield
    finally:
        torch.set_num_threads(old)


def load_reference_callers_on(qu_package, alias: str = "models_on_b200"):
    """Import the reference's UNMODIFIED caller files (vit_quant.py, layers_quant.py, model_utils.py)
    with `.quantization_utils` resolved to `qu_package` (e.g. ivit_b200.quantization_utils): the
    "vit_quant.py runs unchanged on top of it" check.  Returns (vit_quant, model_utils) modules."""
    import importlib

    if alias + ".vit_quant" in sys.modules:
        return sys.modules[alias + ".vit_quant"], sys.modules[alias + ".model_utils"]
    if not reference_available():
        raise RuntimeError(f"reference not mounted at {REFERENCE_ROOT}")
    _install_stubs()
    pkg = types.ModuleType(alias)
    pkg.__path__ = [os.path.join(REFERENCE_ROOT, "models")]      # a package without running models/__init__.py
    pkg.__package__ = alias
    sys.modules[alias] = pkg
    sys.modules[alias + ".quantization_utils"] = qu_package
    vq = importlib.import_module(alias + ".vit_quant")
    mu = importlib.import_module(alias + ".model_utils")
    return vq, mu


def load_reference_swin_on(qu_package, alias: str = "models_on_b200"):
    """The reference's UNMODIFIED swin_quant.py on `qu_package`: needs only the tkinter stub -- the package
    itself exports `IntLayerNorm / IntSoftmax / IntGELU` and accepts `QuantLinear(bias=False)`."""
    import importlib

    load_reference_callers_on(qu_package, alias)
    return importlib.import_module(alias + ".swin_quant")


_SWIN = None


def load_reference_swin():
    """The reference's `models.swin_quant`, importable only with the three fixes of SURVEY.md section 0-2
    (none of them edits the source): the tkinter stub, the `IntLayerNorm / IntSoftmax / IntGELU` names
    that `swin_quant.py:11` imports but `quantization_utils/__init__.py` never exported (= the IVIT
    classes: the only ones whose constructors match the call sites), and -- applied by the caller to
    each built model, see `give_biasless_linears_a_zero_bias` -- a bias for `PatchMerging.reduction`."""
    global _SWIN
    if _SWIN is not None:
        return _SWIN
    import importlib

    ref = load_reference()
    qu = sys.modules[ref.__name__ + ".quantization_utils"]
    qu.IntLayerNorm, qu.IntSoftmax, qu.IntGELU = qu.IVITIntLayerNorm, qu.IVITIntSoftmax, qu.IVITIntGELU
    sys.path.insert(0, REFERENCE_ROOT)
    try:
        _SWIN = importlib.import_module(ref.__name__ + ".swin_quant")
    finally:
        sys.path.remove(REFERENCE_ROOT)
    return _SWIN


def give_biasless_linears_a_zero_bias(model) -> int:
    """`QuantLinear(..., bias=False)` (swin_quant.py:321-325) crashes in the reference's forward
    (quant_modules.py:219-220 quantises `self.bias` unconditionally).  A zero bias is arithmetically
    the upstream behaviour (F.linear without bias): bias_integer = 0 and x + 0.0 == x."""
    import torch

    n = 0
    for m in model.modules():
        if type(m).__name__ == "QuantLinear" and m.bias is None:
            m.bias = torch.nn.Parameter(torch.zeros(m.out_features))
            m.register_buffer("bias_integer", torch.zeros(m.out_features))
            n += 1
    return n
