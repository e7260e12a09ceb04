"""Pins the numpy Swin oracle (oracle/swin_oracle.py) against the committed outputs of the UNMODIFIED
reference swin_quant.py (tests/golden/swin_*.npz, made by oracle/make_golden.py through the shimmed
import of SURVEY.md section 0-2).  Runs anywhere.  Bar: codes and scales bit-equal at every hooked module
boundary (106 / 298 of them), logits equal to 1e-6 relative with identical argmax."""
import hashlib

import numpy as np
import pytest

from ivit_b200.synthetic import load_swin_case, synth_images
from oracle import swin_oracle as SO

f32 = np.float32


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _codes(y, s):
    return np.rint(np.asarray(y, dtype=f32) / np.asarray(s, dtype=f32)).astype(np.int64)


@pytest.mark.parametrize("case", ["swin_toy", "swin_tiny", "swin_toy_b", "swin_toy_w4"])
def test_swin_oracle_matches_reference_fixture(case):
    cfg, state, fix = load_swin_case(case)
    x = synth_images(2, cfg, seed=0).numpy()
    o = SO.SwinOracle(cfg, state)
    logits = o.forward(x, record=True)
    names = list(fix["tap_names"])
    code_sha, scale_sha = dict(zip(names, fix["tap_code_sha"])), dict(zip(names, fix["tap_scale_sha"]))
    assert sorted(o.taps) == sorted(names)
    for k, (y, s) in o.taps.items():
        assert _sha(np.asarray(s, dtype=f32).reshape(np.asarray(fix_shape(k, s)).shape)) == scale_sha[k] or \
            _sha(np.asarray(s, dtype=f32)) == scale_sha[k], f"{case}: scale of {k}"
        assert _sha(_codes(y, s)) == code_sha[k], f"{case}: codes of {k}"
    ref = fix["logits"]
    assert np.array_equal(ref.argmax(1), logits.argmax(1))
    assert np.allclose(logits, ref, rtol=1e-6, atol=1e-6 * np.abs(ref).max())
    for k in fix.files:
        if k.startswith("codes/"):
            name = k[len("codes/"):]
            assert np.array_equal(_codes(*o.taps[name]), fix[k].astype(np.int64)), name


def fix_shape(k, s):
    return s


def test_avgpool_and_strided_sum_orders_are_torch_orders():
    """The two summation orders the Swin path adds: AdaptiveAvgPool1d over tokens (swin_quant.py:554) and the
    LayerNorm mean over a channel-strided tensor (stage 0), both against torch on this host."""
    import torch
    from oracle import ivit_oracle as O
    g = torch.Generator().manual_seed(0)
    for L, C in ((49, 768), (49, 192), (196, 96), (3136, 96)):
        x = torch.randint(-128, 128, (2, L, C), generator=g).float() * 0.0371
        want = torch.nn.AdaptiveAvgPool1d(1)(x.transpose(1, 2))[:, :, 0].numpy()
        assert np.array_equal(SO.avgpool_tokens(x.numpy()), want), (L, C)
    for L, C, Bn in ((196, 96, 2), (3136, 96, 2), (49, 128, 2), (196, 96, 1), (16, 96, 1), (8, 96, 3), (64, 96, 1), (100, 192, 5)):
        q = torch.randint(-32768, 32768, (Bn, C, L), generator=g).float() * 3.1e-4
        xs = q.transpose(1, 2) / torch.tensor(3.1e-4)              # [B, L, C] with channel stride L
        assert not xs.is_contiguous()
        want = xs.sum(dim=2, keepdim=True).numpy()
        assert np.array_equal(O.torch_outer_sum_lastdim(xs.numpy()), want), (L, C, Bn)
        assert np.array_equal((O.torch_outer_sum_lastdim(xs.numpy()) / f32(C)).astype(f32),
                              xs.mean(dim=2, keepdim=True).numpy())
