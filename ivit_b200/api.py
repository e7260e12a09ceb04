"""Public inference call: host images in, host logits out.

``InferenceSession`` owns one ``FrozenVitPlan`` (or ``FrozenSwinPlan``) on one GPU and what a serving loop needs around
it: CUDA-graph replays of the launch sequence, a copy stream, and rotating staging slots so that
the host->device copy of the next chunk (or the next call) overlaps the forward in flight.  This is the
call ``bench.py`` times as ``e2e`` and the one ``scripts/inference.py:248`` (``model(imgs)`` per
batch) maps to.  No CPU path: construction fails without a CUDA device and the in-tree library.
"""
from __future__ import annotations

import torch

from .plan import FrozenVitPlan
from .swin_plan import FrozenSwinPlan
from .synthetic import SwinConfig, VitConfig


class InferenceSession:
    def __init__(self, cfg: "VitConfig | SwinConfig", state: dict, batch: int, device="cuda", chunks: int = 1, slots: int = 2,
                 uint8_input=None):
        """batch images per call, processed as `chunks` equal forwards; `slots` staging slots (each a
        captured graph with its own input buffer) rotate across chunks AND calls, so with slots >= 2
        the host->device copy of the next chunk / next call overlaps the forward in flight."""
        """uint8_input=(mean, std): the session takes uint8 images [batch,C,H,W] (0..255) and applies ToTensor +
        Normalize(mean, std) + the input QuantAct on the device as one code table: a quarter of the float32 input bytes
        over the host->device link, bit-identical logits to feeding the normalised float32 tensor (ViT plans)."""
        if batch % chunks != 0:
            raise ValueError("batch must be divisible by chunks")
        if slots < 1:
            raise ValueError("slots must be >= 1")
        self.plan = FrozenSwinPlan(cfg, state, device) if isinstance(cfg, SwinConfig) else FrozenVitPlan(cfg, state, device)
        self.device = self.plan.device
        self.batch, self.chunks, self.cb = batch, chunks, batch // chunks
        self.compute = torch.cuda.Stream(device=self.device)
        self.copy = torch.cuda.Stream(device=self.device)
        self.readback = torch.cuda.Stream(device=self.device)      # logits device->host, off the compute stream
        self.in_dtype = torch.float32
        if uint8_input is not None:
            if not isinstance(self.plan, FrozenVitPlan):
                raise NotImplementedError("uint8 input: ViT / DeiT plans")
            self.plan.set_uint8_input(*uint8_input)
            self.in_dtype = torch.uint8
        with torch.cuda.device(self.device):
            self.slots = [self.plan.capture(self.cb, slot=s, dtype=self.in_dtype) for s in range(slots)]
            self.ready = [torch.cuda.Event() for _ in self.slots]
            self.consumed = [torch.cuda.Event() for _ in self.slots]
            self.read = [torch.cuda.Event() for _ in self.slots]     # the slot's logits have been copied out
        # two result buffers: the logits of call k stay valid while call k+1 is in flight
        self.out_hosts = [torch.empty((batch, cfg.num_classes), dtype=torch.float32).pin_memory() for _ in range(2)]
        self.h2d_bytes = batch * cfg.in_chans * cfg.img_size * cfg.img_size * (4 if uint8_input is None else 1)
        self.d2h_bytes = batch * cfg.num_classes * 4
        self.launches_per_infer = chunks * self.slots[0].launches
        self._seq = 0          # chunks enqueued so far (slot rotation)
        self._calls = 0

    def infer_async(self, images: torch.Tensor) -> torch.Tensor:
        """Enqueue one batch.  images: float32 [batch,3,H,W] in HOST memory (pinned for an
        asynchronous copy; the caller keeps it unchanged until the copy has run).  Returns the pinned
        host logits buffer of this call; valid after ``synchronize()``, reused two calls later."""
        if images.is_cuda:
            raise ValueError("infer() takes host images; use FrozenVitPlan.forward for device tensors")
        if images.shape[0] != self.batch:
            raise ValueError(f"session was built for batch {self.batch}")
        if images.dtype != self.in_dtype:
            raise ValueError(f"session takes {self.in_dtype} images")
        out_host = self.out_hosts[self._calls % 2]
        self._calls += 1
        ns = len(self.slots)
        for c in range(self.chunks):
            s = self._seq % ns
            self._seq += 1
            g = self.slots[s]
            lo, hi = c * self.cb, (c + 1) * self.cb
            with torch.cuda.stream(self.copy):
                self.copy.wait_event(self.consumed[s])          # the slot's previous forward has read it
                g.static_in.copy_(images[lo:hi], non_blocking=True)
                self.ready[s].record(self.copy)
            with torch.cuda.stream(self.compute):
                self.compute.wait_event(self.ready[s])
                self.compute.wait_event(self.read[s])           # the slot's previous logits are out (no-op at first)
                g.replay()
                self.consumed[s].record(self.compute)
            with torch.cuda.stream(self.readback):               # the next forward does not queue behind this copy
                self.readback.wait_event(self.consumed[s])
                out_host[lo:hi].copy_(g.static_out, non_blocking=True)
                self.read[s].record(self.readback)
        return out_host

    def join(self):
        """Make the compute stream wait for every readback enqueued so far (an event recorded on the compute stream
        afterwards then covers the device->host copies too: bench.py's end-to-end timing)."""
        self.compute.wait_stream(self.readback)

    def synchronize(self):
        self.compute.synchronize()
        self.readback.synchronize()

    def infer(self, images: torch.Tensor) -> torch.Tensor:
        out = self.infer_async(images)
        self.synchronize()
        return out


def swin_config_from_model(model) -> SwinConfig:
    """SwinConfig of a swin_quant.SwinTransformer (or any caller with the same attribute names) built on
    the quantization_utils module API.  The layer types and bit-widths are the ones swin_quant.py
    hard-codes (IntLayerNorm / IntSoftmax / IntGELU = the I-ViT classes; 16-bit residual stream)."""
    stages = list(model.layers)
    b0 = stages[0].blocks[0]
    for st in stages:
        for blk in st.blocks:
            if getattr(blk.norm1, "_kind", 0) != 0 or getattr(blk.attn.log_int_softmax, "_kind", 0) != 0 or \
                    getattr(blk.mlp.act, "_kind", 0) != 0:
                raise NotImplementedError("FrozenSwinPlan covers the I-ViT layer types swin_quant.py is written for")
            if (blk.qact2.activation_bit, blk.qact4.activation_bit, blk.attn.qact4.activation_bit) != (16, 16, 16):
                raise NotImplementedError("FrozenSwinPlan: the 16-bit residual stream of swin_quant.py:110,214,222")
    pe = model.patch_embed
    ws = max(blk.window_size for st in stages for blk in st.blocks)
    return SwinConfig(name="custom", embed_dim=pe.proj.out_channels, depths=tuple(len(st.blocks) for st in stages),
                      num_heads=tuple(st.blocks[0].attn.num_heads for st in stages), window_size=int(ws),
                      mlp_ratio=b0.mlp.fc1.out_features // b0.mlp.fc1.in_features, img_size=int(pe.img_size[0]),
                      patch_size=int(pe.proj.kernel_size[0]), in_chans=pe.proj.in_channels,
                      num_classes=model.head.out_features)


def config_from_model(model):
    """VitConfig (or SwinConfig) of a model built on the quantization_utils module API (the reference's
    vit_quant.py / swin_quant.py classes or any caller with the same attribute names)."""
    if hasattr(model, "layers") and not hasattr(model, "blocks"):
        return swin_config_from_model(model)
    blk = model.blocks[0]
    C = model.head.in_features
    P = model.patch_embed.proj.kernel_size[0]
    tokens = model.pos_embed.shape[1]
    side = int(round((tokens - 1) ** 0.5))

    def kind(mod):
        name = "ivit" if getattr(mod, "_kind", 0) == 0 else "ibert"
        return name + ("_use-int-sqrt_true" if getattr(mod, "use_int_sqrt", False) else "")
    bw = (model.patch_embed.qact.activation_bit, model.qact_pos.activation_bit, model.qact1.activation_bit,
          blk.attn.qact3.activation_bit, blk.attn.int_softmax.output_bit, blk.mlp.qact2.activation_bit,
          blk.qact2.activation_bit, blk.qact4.activation_bit)
    return VitConfig(name="custom", embed_dim=C, depth=len(model.blocks), num_heads=blk.attn.num_heads,
                     mlp_ratio=blk.mlp.fc1.out_features // C, img_size=side * P, patch_size=P,
                     in_chans=model.patch_embed.proj.in_channels, num_classes=model.head.out_features, bitwidths=bw,
                     gelu_type=kind(blk.mlp.act), softmax_type=kind(blk.attn.int_softmax), layernorm_type=kind(blk.norm1))


def compile_frozen(model, batch: int, device="cuda", chunks: int = 1) -> InferenceSession:
    """Frozen module-API model (after calibration + freeze_model) -> the fused static plan: same
    integer results, ~100 launches per forward instead of one Python call per module."""
    return InferenceSession(config_from_model(model), model.state_dict(), batch, device, chunks)
