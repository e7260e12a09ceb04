"""Build-container only: the numpy oracle against the UNMODIFIED reference, imported through
oracle/ref_shim.py.  Skipped wherever /root/reference is absent (the GPU box)."""
import numpy as np
import pytest
import torch

from ivit_b200.synthetic import load_case, synth_images
from oracle import ivit_oracle as O
from oracle import vit_oracle as VO

pytestmark = pytest.mark.reference
f32 = np.float32


@pytest.mark.parametrize("n", [1, 8, 31, 32, 49, 64, 192, 197, 384, 768, 1000, 1536, 3072, 8200])
def test_torch_sum_order(n):
    # (rows of 2..7 elements take a different ATen loop; no row on this path is that short)
    g = torch.Generator().manual_seed(n)
    x = (torch.randn(37, n, generator=g) * 3e6).float()       # partial sums far beyond 2^24
    assert np.array_equal(O.torch_sum_lastdim(x.numpy()), torch.sum(x, dim=-1, keepdim=True).numpy())
    x = torch.round(torch.rand(11, n, generator=g) * 5e8).float()
    assert np.array_equal(O.torch_sum_lastdim(x.numpy()), torch.sum(x, dim=-1, keepdim=True).numpy())


def test_batch_frexp_matches_reference():
    from oracle.ref_shim import load_reference
    import sys
    ref = load_reference()
    qu = sys.modules[ref.__name__ + ".quantization_utils.quant_utils"]
    rng = np.random.default_rng(0)
    r = np.concatenate([rng.random(500) * 10.0 ** rng.integers(-6, 6, 500), -rng.random(20), [0.5, 1.0, 2.0 ** -20]])
    m_ref, e_ref = qu.batch_frexp(torch.from_numpy(r))
    m, e = O.batch_frexp(r)
    assert np.array_equal(m_ref.numpy().astype(np.int64), m)
    assert np.array_equal(e_ref.numpy().astype(np.float64), e)


@pytest.mark.parametrize("case,batch", [("toy_ivit", 3), ("toy_mixed", 3), ("deit_small_ivit", 2)])
def test_vit_oracle_vs_live_reference(case, batch):
    """Fresh inputs (not the fixture's seed): every module boundary, codes and scales bit-exact."""
    from oracle import ref_harness as RH
    cfg, state, _ = load_case(case)
    weights = {k: v for k, v in state.items() if not k.endswith((".x_min", ".x_max", ".shift"))}
    model = RH.build_reference_vit(cfg, weights)
    RH.calibrate_and_freeze(model, synth_images(4, cfg, seed=1))
    for k, v in RH.act_ranges(model).items():                   # same calibration as the fixture
        assert np.array_equal(v, state[k].reshape(-1)), k
    x = synth_images(batch, cfg, seed=5)
    logits_ref, taps = RH.run_with_taps(model, x)
    o = VO.VitOracle(cfg, state)
    logits = o.forward(x.numpy(), record=True)
    for k, (y, s) in o.taps.items():
        ys, ss = taps[k]
        assert np.array_equal(np.asarray(s, dtype=f32).reshape(-1), ss.reshape(-1)), k
        assert np.array_equal(VO.codes_of((y, s)), VO.codes_of((ys, ss))), k
    assert np.array_equal(logits.argmax(1), logits_ref.argmax(1))
    assert np.allclose(logits, logits_ref, rtol=1e-6, atol=1e-6 * np.abs(logits_ref).max())
