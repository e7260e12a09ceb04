"""The module API mirror (ivit_b200.quantization_utils) under the reference's own, UNMODIFIED caller
files (models/vit_quant.py, layers_quant.py, model_utils.py) -- build container only.

On CPU only the float-simulated calibration mode can run (the frozen integer path is CUDA-only and
must refuse CPU tensors).  The calibration forward has to reproduce the reference's x_min / x_max
buffers exactly: they define every frozen scale."""
import numpy as np
import pytest
import torch

import ivit_b200.quantization_utils as qu
from ivit_b200 import _cabi
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
    pg = qu.get_gelu("ppoly_deg_2_seg_8")()
    assert isinstance(pg, qu.PPolyIntGELU) and (pg.segments, pg.degree, pg.N, pg.backend) == (8, 2, 22, "ibert")
    ps = qu.get_softmax("ppoly_exp-bits_12")(8)
    assert isinstance(ps, qu.PPolyIntSoftmax) and (ps.exp_bitwidth, ps.N, ps.segments) == (12, 28, 16)
    assert qu.get_gelu("float") is qu.FloatGELU and qu.get_layernorm("FLOAT") is qu.FloatLayerNorm
    assert qu.get_softmax("float_bitwidth_16")().bitwidth == 16
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
@pytest.mark.parametrize("case", ["toy_ivit", "toy_ibert", "toy_mixed", "deit_tiny_ivit", "toy_ppoly", "toy_ppoly_args"])
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
    # ppoly layers: the segment fit of the calibration pass leaves the reference's integer coefficients behind
    nfix = 0
    for k in fix.files:
        if k.startswith("fixed/"):
            name = k[len("fixed/"):]
            assert np.array_equal(sd[name].numpy(), fix[k]), name
            nfix += 1
    assert nfix == (5 * cfg.depth if "ppoly" in case else 0)
    mu.freeze_model(model)
    assert model.qact_input.running_stat is False
    with pytest.raises(IvitKernelError):                    # frozen = integer kernels = CUDA only
        model(synth_images(1, cfg, seed=0))


@pytest.mark.reference
def test_reference_swin_quant_runs_unchanged_and_calibrates_identically():
    """north_star: swin_quant.py runs unchanged on this package (the reference's own
    quantization_utils cannot run it -- missing Int* exports, bias=False crash).  Same weights, same
    calibration batch as oracle/make_golden.py: every calibrated range equals the (shimmed) reference's,
    i.e. the float-simulated path reproduces it through roll / window partition / bias gather / mask /
    PatchMerging / average pool."""
    from functools import partial
    from ivit_b200.synthetic import load_swin_case
    from oracle.ref_shim import load_reference_swin_on
    sw = load_reference_swin_on(qu)
    cfg, state, fix = load_swin_case("swin_toy")
    model = sw.SwinTransformer(img_size=cfg.img_size, patch_size=cfg.patch_size, in_chans=cfg.in_chans,
                               num_classes=cfg.num_classes, embed_dim=cfg.embed_dim, depths=cfg.depths,
                               num_heads=cfg.num_heads, window_size=cfg.window_size, mlp_ratio=float(cfg.mlp_ratio),
                               norm_layer=partial(qu.IntLayerNorm, eps=1e-6))
    assert isinstance(model.layers[0].blocks[0].attn.qkv, qu.QuantLinear)
    assert model.layers[0].downsample.reduction.bias is None
    weights = {k: torch.as_tensor(v) for k, v in state.items() if not k.endswith((".x_min", ".x_max"))}
    missing, unexpected = model.load_state_dict(weights, strict=False)
    assert not unexpected
    assert not [k for k in missing if k.endswith((".weight", ".bias", "relative_position_bias_table"))], missing
    model.eval()
    with torch.no_grad():
        logits = model(synth_images(4, cfg, seed=1))
    assert logits.shape == (4, cfg.num_classes) and torch.isfinite(logits).all()
    sd = model.state_dict()
    n = 0
    for k in fix.files:
        if k.startswith("range/"):
            name = k[len("range/"):]
            assert np.array_equal(sd[name].reshape(-1).numpy().astype(f32), fix[k].astype(f32)), name
            n += 1
    assert n >= 80
    # the calibrated, frozen model maps onto the fused static plan (compile_frozen's host half)
    from ivit_b200.api import config_from_model
    from ivit_b200.swin_plan import FrozenSwinPlan
    got = config_from_model(model)
    import dataclasses
    assert dataclasses.replace(got, name=cfg.name) == cfg
    plan = FrozenSwinPlan(got, model.state_dict(), host_only=True)
    want = FrozenSwinPlan(cfg, state, host_only=True)
    assert plan.pool_mult == want.pool_mult and plan.stages[1]["blocks"][1]["proj"].mult2 == \
        want.stages[1]["blocks"][1]["proj"].mult2


def _expect_swin_toy_launches(calls, B):
    """swin_toy: 14x14 tokens, 2 + 2 blocks, 7x7 windows (block 1 of stage 0 shifted and masked), one PatchMerging."""
    from collections import Counter
    n = Counter(c.split(":")[0] for c in calls)
    assert n["window_attention"] == 4 and n["gelu"] == 4 and n["avgpool"] == 1 and n["quantize_patches"] == 1
    assert n["ln_strided"] == 5 and n["ln"] == 6          # stage 0 (and the stem) channel-strided in the reference: H11
    wa = [c for c in calls if c.startswith("window_attention")]
    assert wa[0] == f"window_attention:B={4 * B}:res=7:ws=7:shift=0:nH=3:windowed=1"      # every window on its own
    assert wa[1] == f"window_attention:B={B}:res=14:ws=7:shift=3:nH=3:windowed=1"         # mask checked, geometry kept
    assert wa[2] == wa[3] == f"window_attention:B={B}:res=7:ws=7:shift=0:nH=6:windowed=1"
    assert f"gemm:{_cabi.EPI_LOGITS}" == calls[-1]


def test_frozen_swin_forward_walks_the_carrier(monkeypatch):
    """Row a16, host logic (no GPU, nothing computed: tests/kernel_stubs.py): a frozen swin_quant-style forward stays
    on integer carriers through torch.roll, window_partition / window_reverse, the bias gather, the mask add,
    PatchMerging and the average pool, and maps to the expected launches."""
    import kernel_stubs
    from ivit_b200 import lifecycle
    from ivit_b200.synthetic import load_swin_case
    from swin_caller import SwinTransformer
    cfg, state, _ = load_swin_case("swin_toy")
    model = SwinTransformer(cfg)
    lifecycle.load_checkpoint(model, {"model": {k: torch.as_tensor(v) for k, v in state.items()}})
    lifecycle.freeze_model(model, fused_plan=False)
    calls = kernel_stubs.install(monkeypatch)
    with torch.no_grad():
        out = model(synth_images(2, cfg, seed=0))
        assert isinstance(out, qu.QTensor) and tuple(out.shape) == (2, cfg.num_classes)
        assert out.dequantize().shape == (2, cfg.num_classes)
    _expect_swin_toy_launches(calls, 2)
    # a mask that is not the shifted-window mask is refused, not silently replaced
    model.layers[0].blocks[1].attn_mask = model.layers[0].blocks[1].attn_mask * 0.5
    with torch.no_grad(), pytest.raises(NotImplementedError):
        model(synth_images(2, cfg, seed=0))


@pytest.mark.reference
def test_reference_swin_quant_frozen_forward_walks_the_carrier(monkeypatch):
    """The same walk under the reference's UNMODIFIED swin_quant.py, frozen by the reference's own freeze_model."""
    import kernel_stubs
    from functools import partial
    from ivit_b200 import lifecycle
    from ivit_b200.synthetic import load_swin_case
    from oracle.ref_shim import load_reference_swin_on
    sw = load_reference_swin_on(qu)
    cfg, state, _ = load_swin_case("swin_toy")
    model = sw.SwinTransformer(img_size=cfg.img_size, patch_size=cfg.patch_size, in_chans=cfg.in_chans,
                               num_classes=cfg.num_classes, embed_dim=cfg.embed_dim, depths=cfg.depths,
                               num_heads=cfg.num_heads, window_size=cfg.window_size, mlp_ratio=float(cfg.mlp_ratio),
                               norm_layer=partial(qu.IntLayerNorm, eps=1e-6))
    lifecycle.load_checkpoint(model, {"model": {k: torch.as_tensor(v) for k, v in state.items()}})
    model.eval()
    for m in model.modules():                   # model_utils.py:13-15
        if hasattr(m, "fix") and callable(m.fix):
            m.fix()
    calls = kernel_stubs.install(monkeypatch)
    with torch.no_grad():
        out = model(synth_images(3, cfg, seed=0))
    assert isinstance(out, qu.QTensor) and tuple(out.shape) == (3, cfg.num_classes)
    out.dequantize()
    _expect_swin_toy_launches(calls, 3)


def test_swin_caller_calibrates_to_the_reference_ranges():
    """tests/swin_caller.py (the Swin caller the GPU tests use: the reference file cannot travel) reproduces the
    reference's calibration on the same weights and batch: every x_min / x_max equals the fixture's, which was
    produced by the unmodified swin_quant.py.  Runs anywhere."""
    from ivit_b200.synthetic import load_swin_case
    from swin_caller import SwinTransformer
    cfg, state, fix = load_swin_case("swin_toy")
    model = SwinTransformer(cfg)
    weights = {k: torch.as_tensor(v) for k, v in state.items() if not k.endswith((".x_min", ".x_max"))}
    missing, unexpected = model.load_state_dict(weights, strict=False)
    assert not unexpected, unexpected
    assert not [k for k in missing if k.endswith((".weight", ".bias", "relative_position_bias_table"))], missing
    model.eval()
    with torch.no_grad():
        model(synth_images(4, cfg, seed=1))
    sd = model.state_dict()
    n = 0
    for k in fix.files:
        if k.startswith("range/"):
            name = k[len("range/"):]
            assert np.array_equal(sd[name].reshape(-1).numpy().astype(f32), fix[k].astype(f32)), name
            n += 1
    assert n >= 80
    from ivit_b200.api import config_from_model
    import dataclasses
    assert dataclasses.replace(config_from_model(model), name=cfg.name) == cfg
    # frozen = integer kernels = CUDA only, also for the plan-routed Swin forward; unfreeze restores the module forward
    from ivit_b200 import lifecycle
    lifecycle.freeze_model(model)
    with pytest.raises(IvitKernelError):
        model(synth_images(1, cfg, seed=0))
    lifecycle.unfreeze_model(model)
    with torch.no_grad():
        assert model(synth_images(1, cfg, seed=0)).shape == (1, cfg.num_classes)


def _build_pair(case):
    """The reference's VisionTransformer twice: on its own quantization_utils and on ours, same weights."""
    from oracle.ref_shim import load_reference, load_reference_callers_on
    import sys
    ref = load_reference()
    ref_vq = sys.modules[ref.__name__ + ".vit_quant"]
    vq, mu = load_reference_callers_on(qu)
    cfg, state, fix = load_case(case)
    kw = dict(patch_size=cfg.patch_size, embed_dim=cfg.embed_dim, depth=cfg.depth, num_heads=cfg.num_heads,
              mlp_ratio=cfg.mlp_ratio, qkv_bias=True, gelu_type=cfg.gelu_type, softmax_type=cfg.softmax_type,
              layernorm_type=cfg.layernorm_type)
    weights = {k: torch.as_tensor(v) for k, v in state.items() if not k.endswith((".x_min", ".x_max", ".shift"))}
    models = []
    for mod in (ref_vq, vq):
        m = mod.VisionTransformer(**kw)
        m.load_state_dict(weights, strict=False)
        models.append(m.eval())
    return cfg, models[0], models[1]


@pytest.mark.reference
@pytest.mark.parametrize("case", ["toy_ivit", "toy_ibert", "toy_mixed"])
def test_state_dict_wire_format_equals_reference_after_calibration(case):
    """same checkpoint keys, and after the calibration pass EVERY buffer the TVM
    exporter and a later strict load read -- x_min/x_max, act/fc/conv/norm scaling factors,
    weight_integer, bias_integer -- holds the reference's value bit for bit."""
    from ivit_b200 import lifecycle
    cfg, ref_model, our_model = _build_pair(case)
    batches = [synth_images(4, cfg, seed=1), synth_images(2, cfg, seed=5)]     # two batches: EMA path too
    with torch.no_grad():
        for x in batches:
            ref_model(x)
    assert lifecycle.calibrate_model(our_model, batches) == 2
    ref_sd, our_sd = ref_model.state_dict(), our_model.state_dict()
    assert list(ref_sd.keys()) == list(our_sd.keys())
    for k in ref_sd:
        a, b = ref_sd[k], our_sd[k]
        assert a.shape == b.shape and a.dtype == b.dtype, k
        assert torch.equal(a, b) or (torch.isnan(a) == torch.isnan(b)).all() and torch.equal(
            torch.nan_to_num(a), torch.nan_to_num(b)), k
    params = lifecycle.tvm_params(our_model)
    assert params["patch_embed.proj.weight_integer"].dtype == np.int8
    assert params["head.bias_integer"].dtype == np.int32 and "blocks.0.norm1.norm_scaling_factor" in params or \
        cfg.layernorm_type != "ivit"
    # strict reload of our own state_dict (inference.py --strict-load path) and the 0-d buffer fix-up
    cfg2, _, fresh = _build_pair(case)
    sd = {k: (v.reshape(()) if v.numel() == 1 and k.endswith("act_scaling_factor") else v) for k, v in our_sd.items()}
    missing, unexpected = lifecycle.load_checkpoint(fresh, {"model": sd, "epoch": 3}, strict=True)
    assert not missing and not unexpected
    lifecycle.freeze_model(fresh)
    assert fresh.qact_input.running_stat is False
    lifecycle.unfreeze_model(fresh)
    assert fresh.qact_input.running_stat is True


@pytest.mark.reference
def test_float_registry_entries_equal_reference():
    """SURVEY 8(f)-4, `float` part: same outputs and scales as layer_selection.py's Float* wrappers
    (same torch ops, so bit-equal on the same device), and a whole model built on them calibrates to
    the reference's logits and buffers."""
    from oracle.ref_shim import load_reference
    import sys
    ref = load_reference()
    ref_ls = sys.modules[ref.__name__ + ".quantization_utils.layer_selection"]
    g = torch.Generator().manual_seed(0)
    x = torch.randn(2, 7, 16, generator=g) * 2
    s = torch.tensor(0.031)
    for name, args in (("FloatGELU", ()), ("FloatSoftmax", ()), ("FloatLayerNorm", (16,))):
        a, b = getattr(ref_ls, name)(*args), getattr(qu, name)(*args)
        b.load_state_dict(a.state_dict())
        (ya, sa), (yb, sb) = a(x, s), b(x, s)
        assert torch.equal(ya, yb) and torch.equal(torch.as_tensor(sa), torch.as_tensor(sb)), name
    from oracle.ref_shim import load_reference_callers_on
    ref_vq = sys.modules[ref.__name__ + ".vit_quant"]
    vq, mu = load_reference_callers_on(qu)
    cfg, state, fix = load_case("toy_ivit")
    kw = dict(patch_size=cfg.patch_size, embed_dim=cfg.embed_dim, depth=cfg.depth, num_heads=cfg.num_heads,
              mlp_ratio=cfg.mlp_ratio, qkv_bias=True, gelu_type="float", softmax_type="float", layernorm_type="float")
    torch.manual_seed(3)
    m_ref = ref_vq.VisionTransformer(**kw).eval()
    m_our = vq.VisionTransformer(**kw).eval()
    assert list(m_ref.state_dict().keys()) == list(m_our.state_dict().keys())
    m_our.load_state_dict(m_ref.state_dict())
    xin = synth_images(3, cfg, seed=2)
    with torch.no_grad():
        la, lb = m_ref(xin), m_our(xin)
    assert torch.equal(la, lb)
    for k, v in m_ref.state_dict().items():
        assert torch.equal(v, m_our.state_dict()[k]), k


def test_fused_swin_forward_drops_its_plans_on_load_state_dict():
    """lifecycle._FusedPlanForward keeps FrozenSwinPlans built from the state_dict: loading new weights must start over."""
    from ivit_b200 import lifecycle
    from ivit_b200.synthetic import load_swin_case
    from swin_caller import SwinTransformer
    cfg, state, _ = load_swin_case("swin_toy")
    model = SwinTransformer(cfg)
    ck = {"model": {k: torch.as_tensor(v) for k, v in state.items()}}
    lifecycle.load_checkpoint(model, ck)
    lifecycle.freeze_model(model)
    fwd = model.__dict__["forward"]
    fwd._plans[("cuda", 0)] = object()
    lifecycle.load_checkpoint(model, ck)
    assert fwd._plans == {}
    with pytest.raises(IvitKernelError):
        model(synth_images(1, cfg, seed=0))            # CPU tensor: the frozen forward refuses, no fallback
    lifecycle.unfreeze_model(model)
    assert "forward" not in model.__dict__


@pytest.mark.reference
@pytest.mark.parametrize("momentum", [0.95, -1])
def test_percentile_and_minmax_calibration_modes_equal_reference(momentum):
    """QuantAct's other running-stat modes (quant_modules.py:319-355): percentile ranges (torch.quantile) and the
    running min / max (act_range_momentum = -1) -- every calibrated buffer equals the reference's."""
    from ivit_b200 import lifecycle
    cfg, ref_model, our_model = _build_pair("toy_ivit")
    for model in (ref_model, our_model):
        n = 0
        for m in model.modules():
            if type(m).__name__ == "QuantAct":
                m.percentile = 99.9 if n % 2 == 0 else None          # half of the locations on percentiles
                m.act_range_momentum = momentum
                n += 1
        assert n >= 20
    batches = [synth_images(3, cfg, seed=2), synth_images(2, cfg, seed=6), synth_images(2, cfg, seed=7)]
    with torch.no_grad():
        for x in batches:
            ref_model(x)
    assert lifecycle.calibrate_model(our_model, batches) == 3
    ref_sd, our_sd = ref_model.state_dict(), our_model.state_dict()
    for k in ref_sd:
        if k.endswith((".x_min", ".x_max", "act_scaling_factor")):
            assert torch.equal(ref_sd[k], our_sd[k]), k
