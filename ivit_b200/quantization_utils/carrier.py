"""QTensor: the integer-code carrier the frozen modules exchange on the GPU.

The reference passes ``codes * scale`` as a float32 tensor plus a separate scale, and its callers
(vit_quant.py:61-90, :285-312) only ever reshape / permute / slice / cat that tensor, multiply it by
a python float, or hand it to the next module.  QTensor keeps the int8/int32 codes instead and
tracks those view operations on a ``meta`` tensor (sizes / strides / offset over a canonical
storage), so nothing is computed or copied until a module consumes it.  A QTensor may also be
*pending*: it names the producer (GEMM, LayerNorm, GELU, attention chain, quantiser, cat) whose
kernel is launched -- fused with the consumer's QuantAct -- when the consumer arrives.

It is a duck type, not a torch.Tensor subclass: ``__torch_function__`` routes the torch.* calls the
callers make (torch.cat, torch.flatten, F.dropout in eval, ...); anything unknown dequantises to the
reference's float carrier first, so user code (argmax, loss, hooks) keeps working.

swin_quant.py's glue (:18-50 window_partition / window_reverse, :258-289 torch.roll both ways, :140-147 the
relative-position-bias gather, :149-153 the float mask add, :339-345 PatchMerging's slices + cat, :554 the token
average pool) is carried the same way: views stay views; roll / a permuted ``contiguous().view`` / a tensor-index
gather move the integer codes with torch's copy kernels into a new canonical storage; the mask add and the average
pool stay symbolic until the module that consumes them launches the fused kernel.  ``layout`` remembers what the
reference's float tensor would look like in memory (DESIGN.md hazard H11: torch's summation order over the channel
dimension depends on it).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

VIEW_METHODS = ("reshape", "view", "permute", "transpose", "flatten", "squeeze", "unsqueeze", "expand",
                "contiguous_view")


class Pending:
    """A producer whose kernel has not been launched yet."""
    kind = "?"

    def __init__(self, kind, **kw):
        self.kind = kind
        self.__dict__.update(kw)


class QTensor:
    __slots__ = ("codes", "pending", "meta", "scale", "hscale", "factor", "canon_shape", "device", "layout")

    def __init__(self, codes=None, pending=None, canon_shape=None, scale=None, meta=None, factor=1.0,
                 device=None, hscale=None, layout=0):
        self.codes = codes                 # canonical storage (contiguous), or None while pending
        self.pending = pending
        self.canon_shape = tuple(codes.shape) if codes is not None else tuple(canon_shape)
        self.device = codes.device if codes is not None else device
        self.scale = scale                 # float32 device tensor: value = codes * scale (* factor)
        self.hscale = hscale               # host copy (numpy float32) of `scale`: frozen constants are built from it
        self.factor = factor               # python-float multipliers applied by the caller (attn * self.scale)
        self.meta = meta if meta is not None else torch.empty(self.canon_shape, device="meta", dtype=torch.int8)
        # 0: the reference's float tensor is contiguous along the channel (last) dimension.  L > 0: it is the
        # channel-strided [B, L, C] view of an NCHW conv output (layers_quant.py:198) or an elementwise result of
        # one -- LayerNorm's mean over it follows ATen's outer-reduction order, L tokens per image
        self.layout = layout

    # ------------------------------------------------------------------ shape protocol
    @property
    def shape(self):
        return self.meta.shape

    def size(self, *a):
        return self.meta.size(*a)

    def dim(self):
        return self.meta.dim()

    @property
    def ndim(self):
        return self.meta.dim()

    def numel(self):
        return self.meta.numel()

    @property
    def is_cuda(self):
        return self.device is not None and self.device.type == "cuda"

    def _with_meta(self, meta):
        return QTensor(self.codes, self.pending, self.canon_shape, self.scale, meta, self.factor, self.device,
                       self.hscale, self.layout)

    def is_canonical(self) -> bool:
        """The current view is the whole canonical storage in order (any reshape of it)."""
        return self.meta.is_contiguous() and self.meta.numel() == _prod(self.canon_shape) and \
            self.meta.storage_offset() == 0

    def strided(self, storage: torch.Tensor) -> torch.Tensor:
        """Apply this carrier's view to a tensor laid out like its canonical storage."""
        return torch.as_strided(storage.reshape(-1), tuple(self.meta.shape), tuple(self.meta.stride()),
                                self.meta.storage_offset())

    # ------------------------------------------------------------------ views (no data movement)
    def _view_op(self, name, *a, **k):
        try:
            m = getattr(self.meta, name)(*a, **k)
        except RuntimeError:
            m = None
        if m is None or (name == "reshape" and not _aliases(m, self.meta)):
            # not expressible as a view: materialise codes, do the real op, start a new canonical storage
            # (`view` after the caller's `.contiguous()`, swin_quant.py:30-31: contiguous() is lazy here -> reshape)
            t = getattr(self.materialize(), "reshape" if name == "view" else name)(*a, **k).contiguous()
            return QTensor(t, scale=self.scale, factor=self.factor, hscale=self.hscale)
        return self._with_meta(m)

    def reshape(self, *shape):
        return self._view_op("reshape", *shape)

    def view(self, *shape):
        return self._view_op("view", *shape)

    def permute(self, *dims):
        return self._view_op("permute", *dims)

    def transpose(self, a, b):
        return self._view_op("transpose", a, b)

    def flatten(self, start_dim=0, end_dim=-1):
        return self._view_op("flatten", start_dim, end_dim)

    def squeeze(self, *a):
        return self._view_op("squeeze", *a)

    def unsqueeze(self, d):
        return self._view_op("unsqueeze", d)

    def expand(self, *a):
        return self._view_op("expand", *a)

    def __getitem__(self, idx):
        if isinstance(idx, torch.Tensor) or (isinstance(idx, tuple) and any(isinstance(i, torch.Tensor) for i in idx)):
            # gather by an index tensor (swin_quant.py:142: bias table rows by relative_position_index): new storage
            t = self.materialize()[idx].contiguous()
            return QTensor(t, scale=self.scale, factor=self.factor, hscale=self.hscale)
        return self._with_meta(self.meta[idx])

    def contiguous(self):
        if self.layout and self.codes is not None:
            q = self._with_meta(self.meta)
            q.layout = 0                       # the reference's copy is contiguous from here on
            return q
        return self

    def roll(self, shifts, dims=None):
        """torch.roll of the integer codes (swin_quant.py:258-262, :283-287); the result is a fresh contiguous
        tensor in the reference as well."""
        t = torch.roll(self.materialize(), shifts, dims).contiguous()
        return QTensor(t, scale=self.scale, factor=self.factor, hscale=self.hscale)

    def __add__(self, other):
        p = self.pending if self.codes is None else None
        if p is not None and p.kind == "qattn_bias" and isinstance(other, torch.Tensor) and p.mask is None \
                and not isinstance(other, QTensor):
            # attn.view(B, nW, nH, N, N) + mask (swin_quant.py:149-152): kept symbolic, the window-attention kernel
            # regenerates the 0 / -100 mask from the window geometry after checking it against this tensor
            m = self.meta + torch.empty(other.shape, device="meta", dtype=torch.int8)
            if tuple(m.shape) == tuple(self.meta.shape) and self.is_canonical():
                q = self._with_meta(self.meta)
                q.pending = Pending("qattn_bias", **{k: v for k, v in p.__dict__.items() if k != "kind"})
                q.pending.mask = other
                return q
        return self.dequantize() + (other.dequantize() if isinstance(other, QTensor) else other)

    __radd__ = __add__

    def detach(self):
        return self

    def __mul__(self, other):
        if isinstance(other, (int, float)):
            q = self._with_meta(self.meta)
            q.factor = self.factor * float(other)
            return q
        return self.dequantize() * other

    __rmul__ = __mul__

    # ------------------------------------------------------------------ materialisation
    def materialize(self) -> torch.Tensor:
        """Integer codes of the current view (launches the pending producer un-fused if needed)."""
        if self.codes is None:
            from . import frozen
            self.codes = frozen.run_pending(self)
            self.pending = None
        return self.strided(self.codes)

    def dequantize(self) -> torch.Tensor:
        """The reference's float carrier: codes * scale (per-channel scales broadcast on the last dim)."""
        from . import frozen
        return frozen.dequantize(self)

    # ------------------------------------------------------------------ torch.* routing
    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        if func in (F.dropout, torch.dropout) and not kwargs.get("training", args[2] if len(args) > 2 else True):
            return args[0]
        if func is torch.cat:
            from . import frozen
            return frozen.cat(*args, **kwargs)
        if func is torch.flatten:
            return args[0].flatten(*args[1:], **kwargs)
        if func is torch.transpose:
            return args[0].transpose(*args[1:])
        if func is torch.reshape:
            return args[0].reshape(*args[1:])
        if func is torch.permute:
            return args[0].permute(*args[1:])
        if func is torch.roll and isinstance(args[0], QTensor):
            return args[0].roll(*args[1:], **kwargs)
        if func in (F.adaptive_avg_pool1d, torch.adaptive_avg_pool1d) and isinstance(args[0], QTensor):
            from . import frozen
            return frozen.avgpool1d(args[0], *args[1:], **kwargs)
        # anything else sees the float carrier
        conv = lambda a: a.dequantize() if isinstance(a, QTensor) else a
        args = tuple(type(a)(conv(x) for x in a) if isinstance(a, (list, tuple)) else conv(a) for a in args)
        kwargs = {k: conv(v) for k, v in kwargs.items()}
        return func(*args, **kwargs)

    def __getattr__(self, name):
        # user-side tensor methods (argmax, topk, float, cpu, ...) act on the float carrier
        if name.startswith("__"):
            raise AttributeError(name)
        return getattr(self.dequantize(), name)

    def __repr__(self):
        st = "pending:" + self.pending.kind if self.codes is None else str(self.codes.dtype)
        return f"QTensor(shape={tuple(self.shape)}, {st}, factor={self.factor})"


def _prod(shape):
    n = 1
    for s in shape:
        n *= int(s)
    return n


def _aliases(a: torch.Tensor, b: torch.Tensor) -> bool:
    """meta tensors: is `a` (the result of a reshape of `b`) a view rather than a copy?"""
    return a is b or a._is_view()
