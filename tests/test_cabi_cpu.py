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
