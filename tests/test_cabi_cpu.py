"""CPU: the C-ABI library builds for sm_100a, loads without a GPU and exports every symbol the
header declares; compute calls refuse to run without CUDA."""
import os
import re

import pytest
import torch

from ivit_b200 import _cabi, build, ops

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_header_symbol():
    lib = _cabi.load()
    hdr = open(os.path.join(ROOT, "include", "ivit_b200.h")).read()
    declared = set(re.findall(r"^\s*(?:int|const char\*)\s+(ivit_\w+)\s*\(", hdr, flags=re.M))
    assert declared, "no declarations parsed"
    assert declared == set(_cabi.EXPORTS), declared ^ set(_cabi.EXPORTS)
    for name in declared:
        assert hasattr(lib, name)
    assert lib.ivit_abi_version() >= 1


def test_sass_is_blackwell_native():
    """tcgen05.mma -> UTC*MMA, tcgen05.ld -> LDTM, TMA -> UTMALDG in the sm_100a SASS."""
    import shutil
    import subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.isfile(cuobjdump):
        pytest.skip("cuobjdump not available")
    sass = subprocess.run([cuobjdump, "-sass", build.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in sass
    for mnemonic in ("UTCIMMA", "LDTM", "UTMALDG"):
        assert mnemonic in sass, mnemonic
    assert "HMMA" not in sass and "IMMA." not in sass.replace("UTCIMMA.", "")


def test_no_cpu_fallback():
    with pytest.raises(_cabi.IvitKernelError):
        ops.quantize(torch.zeros(4, 4), 1.0)


def _rn32(x):
    """Exact rational -> nearest float32 (ties to even), no double rounding."""
    import numpy as np
    from fractions import Fraction
    c = np.float32(float(x))
    cands = {c, np.nextafter(c, np.float32(np.inf)), np.nextafter(c, np.float32(-np.inf))}
    best = None
    for v in cands:
        d = abs(Fraction(float(v)) - x)
        even = (int(np.float32(v).view(np.uint32)) & 1) == 0
        key = (d, 0 if even else 1)
        if best is None or key < best[0]:
            best = (key, v)
    return best[1]


def test_xeff_two_fma_check_is_the_ieee_quotient():
    """ivit_xeff_fma_ok (host arithmetic inside the library, no GPU): where it says yes, the two-fused-operation
    quotient the 16-bit LayerNorm kernels use equals numpy's IEEE float32 x / s -- re-derived here in exact rational
    arithmetic with one rounding per operation."""
    import numpy as np
    from fractions import Fraction as F
    lib = _cabi.load()
    rng = np.random.default_rng(5)
    scales = [np.float32(v) for v in (1.37e-4, 3.1e-5, 0.0371, 2.0 ** -12, 1.0)] + \
        [np.float32(np.exp(rng.uniform(np.log(1e-6), np.log(0.5)))) for _ in range(6)]
    assert lib.ivit_xeff_fma_ok(0.0, 16) == 0 and lib.ivit_xeff_fma_ok(float("nan"), 16) == 0
    assert lib.ivit_xeff_fma_ok(1.0, 12) == 0
    checked = 0
    for s in scales:
        if not lib.ivit_xeff_fma_ok(float(s), 16):
            continue
        inv = np.float32(1.0) / s
        for q in [-32768, -32767, -1, 0, 1, 3, 32767] + [int(v) for v in rng.integers(-32768, 32768, size=60)]:
            r = np.float32(q) * s
            ref = r / s
            c = r * inv
            e = _rn32(F(float(r)) - F(float(c)) * F(float(s)))
            c2 = _rn32(F(float(e)) * F(float(inv)) + F(float(c)))
            assert np.float32(c2) == ref, (float(s), q)
            checked += 1
    assert checked >= 300
