"""The module API mirror (ivit_b200.quantization_utils) under the reference's own, UNMODIFIED caller
files (models/vit_quant.py, layers_quant.py, model_utils.py) -- build container only.

On CPU only the float-simulated calibration mode can run (the frozen integer path is CUDA-only and
must refuse CPU tensors).  The calibration forward has to reproduce the reference's x_min / x_max
buffers exactly: they define every frozen scale."""
import numpy as np
import pytest
import torch

import ivit_b200.quantization_utils as qu
from ivit_b200._cabi import IvitKernelError
from ivit_b200.synthetic import load_case, synth_images

f32 = np.float32


def test_export_list_matches_reference_plus_swin_aliases():
    ref_exports = ["QuantLinear", "QuantAct", "QuantConv2d", "QuantMatMul", "attach_io_stat_hooks", "save_io_stats_df",
                   "enable_io_stats", "disable_io_stats", "clear_io_stats", "softmax", "IVITIntGELU", "IVITIntSoftmax",
                   "IVITIntLayerNorm", "IBERTIntGELU", "IBERTIntSoftmax", "IBERTIntLayerNorm", "get_gelu",
                   "get_softmax", "get_layernorm"]
    for n in ref_exports + ["IntLayerNorm", "IntSoftmax", "IntGELU"]:
        assert hasattr(qu, n), n
    assert qu.IntGELU is qu.IVITIntGELU


def test_registry_grammar():
    cls = qu.get_layernorm("ibert_use-int-sqrt_true")
    assert issubclass(cls, qu.IBERTIntLayerNorm) and cls(64).use_int_sqrt is True
    assert qu.get_gelu("ivit") is qu.IVITIntGELU and qu.get_softmax("IBERT") is qu.IBERTIntSoftmax
    with pytest.raises(NotImplementedError):
        qu.get_gelu("ppoly_deg_2_seg_16")
    with pytest.raises(KeyError):
        qu.get_gelu("nonsense")


def test_buffer_names_are_the_checkpoint_keys():
    assert set(dict(qu.QuantAct().named_buffers())) == {"x_min", "x_max", "act_scaling_factor"}
    assert set(dict(qu.QuantLinear(8, 4).named_buffers())) == {"fc_scaling_factor", "weight_integer", "bias_integer"}
    assert set(dict(qu.QuantConv2d(3, 4, 2, 2).named_buffers())) == {"conv_scaling_factor", "weight_integer",
                                                                      "bias_integer"}
    assert set(dict(qu.IVITIntLayerNorm(8).named_buffers())) == {"norm_scaling_factor", "bias_integer"}
    assert set(dict(qu.IBERTIntLayerNorm(8).named_buffers())) == {"shift"}
    assert set(dict(qu.IBERTIntSoftmax(8).named_buffers())) == {"act.x_min", "act.x_max", "act.act_scaling_factor"}
    assert qu.QuantLinear(8, 4, bias=False).bias is None          # bias=False is legal here


def test_frozen_path_refuses_cpu():
    act = qu.QuantAct()
    act.x_max.fill_(1.0)
    act.fix()
    with pytest.raises(IvitKernelError):
        act(torch.randn(2, 3, 8, 8))


@pytest.mark.reference
@pytest.mark.parametrize("case", ["toy_ivit", "toy_ibert", "toy_mixed", "deit_tiny_ivit"])
def test_reference_vit_quant_runs_unchanged_and_calibrates_identically(case):
    from oracle.ref_shim import load_reference_callers_on
    vq, mu = load_reference_callers_on(qu)
    cfg, state, fix = load_case(case)
    kw = dict(gelu_type=cfg.gelu_type, softmax_type=cfg.softmax_type, layernorm_type=cfg.layernorm_type)
    if cfg.name == "toy":
        model = vq.VisionTransformer(patch_size=cfg.patch_size, embed_dim=cfg.embed_dim, depth=cfg.depth,
                                     num_heads=cfg.num_heads, mlp_ratio=cfg.mlp_ratio, qkv_bias=True, **kw)
    else:
        model = vq.deit_tiny_patch16_224(pretrained=False, **kw)
    assert isinstance(model.qact_input, qu.QuantAct) and isinstance(model.blocks[0].attn.qkv, qu.QuantLinear)
    weights = {k: torch.as_tensor(v) for k, v in state.items() if not k.endswith((".x_min", ".x_max", ".shift"))}
    missing, unexpected = model.load_state_dict(weights, strict=False)
    assert not unexpected
    model.eval()
    with torch.no_grad():
        logits = model(synth_images(4, cfg, seed=1))       # the calibration pass of oracle/make_golden.py
    assert logits.shape == (4, 1000) and torch.isfinite(logits).all()
    sd = model.state_dict()
    n = 0
    for k in fix.files:
        if k.startswith("range/"):
            name = k[len("range/"):]
            got = sd[name].reshape(-1).numpy().astype(f32)
            assert np.array_equal(got, fix[k].astype(f32)), name
            n += 1
    assert n >= 40
    mu.freeze_model(model)
    assert model.qact_input.running_stat is False
    with pytest.raises(IvitKernelError):                    # frozen = integer kernels = CUDA only
        model(synth_images(1, cfg, seed=0))
