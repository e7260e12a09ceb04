"""GPU: randomized self-consistency of the round-2 kernels (restructured LayerNorms, thread-per-row window attention)
against the kernels they replaced -- tools/stress_variants.py over random scales, widths, spreads and window geometries."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_kernel_families_agree_on_random_cases():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "stress_variants.py"), "80"], capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "mismatching cases: 0" in r.stdout
