"""Host-side owner of one ``isg_plan`` (include/isg.h): device memory and streams come from
torch, every computation happens inside libisg.so."""
import ctypes as C

import torch

from . import _lib

_MODES = {"fp32": _lib.MODE_FP32, "bf16": _lib.MODE_BF16, "bf16x3": _lib.MODE_BF16X3}


def state_names(in_channel):
    l = _lib.lib()
    return [l.isg_state_name(in_channel, i).decode() for i in range(l.isg_state_count(in_channel))]


class SegmentEngine:
    """Plan for ``Segment(in_channel)`` at a fixed ``H x W`` on one CUDA device."""

    def __init__(self, device, in_channel, H, W, max_batch, precision="fp32"):
        if precision not in _MODES:
            raise ValueError(f"precision must be one of {sorted(_MODES)}, got {precision!r}")
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _lib.IsgError("SegmentEngine needs a CUDA device; there is no CPU path")
        self.in_channel, self.H, self.W = in_channel, H, W
        self.max_batch, self.precision = max_batch, precision
        self.weights_epoch = None
        self._l = _lib.lib()
        h = C.c_void_p()
        dev = self.device.index if self.device.index is not None else torch.cuda.current_device()
        _lib.check(self._l.isg_plan_create(dev, in_channel, H, W, max_batch, _MODES[precision], C.byref(h)))
        self._h = h
        self._names = state_names(in_channel)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._l.isg_plan_destroy(h)
            except Exception:
                pass

    # ---- weights ----
    def load_state_dict(self, sd):
        """Pack a reference-format ``state_dict`` (train_instance.py:325) into the plan."""
        missing = [k for k in self._names if k not in sd]
        if missing:
            raise KeyError(f"state_dict is missing {len(missing)} keys, e.g. {missing[:3]}")
        keep, ptrs, numels = [], (C.c_void_p * len(self._names))(), (C.c_int64 * len(self._names))()
        for i, k in enumerate(self._names):
            t = sd[k].detach()
            t = t.to("cpu", torch.int64 if k.endswith("num_batches_tracked") else torch.float32).contiguous()
            keep.append(t)
            ptrs[i], numels[i] = t.data_ptr(), t.numel()
        _lib.check(self._l.isg_plan_load_weights(self._h, ptrs, numels, len(self._names)))

    # ---- forward ----
    def forward(self, x, out=None):
        if not (x.is_cuda and x.dtype == torch.float32 and x.dim() == 4):
            raise TypeError("SegmentEngine.forward: x must be a CUDA float32 [B,C,H,W] tensor")
        if tuple(x.shape[1:]) != (self.in_channel, self.H, self.W):
            raise ValueError(f"plan is for [B,{self.in_channel},{self.H},{self.W}], got {tuple(x.shape)}")
        x = x.contiguous()
        B = x.shape[0]
        if out is None:
            out = torch.empty((B, 1, self.H, self.W), device=x.device, dtype=torch.float32)
        with torch.cuda.device(x.device):
            st = torch.cuda.current_stream().cuda_stream
            _lib.check(self._l.isg_forward(self._h, x.data_ptr(), B, out.data_ptr(), st))
        return out

    def forward_u8(self, x, mean, std, out=None):
        """``x``: CUDA uint8 [B,H,W,C] (what an image decoder produces); ToTensor + Normalize(mean, std)
        of the reference's data path (train_instance.py:80-85) run inside the first kernel."""
        if not (x.is_cuda and x.dtype == torch.uint8 and x.dim() == 4):
            raise TypeError("SegmentEngine.forward_u8: x must be a CUDA uint8 [B,H,W,C] tensor")
        if tuple(x.shape[1:]) != (self.H, self.W, self.in_channel):
            raise ValueError(f"plan is for [B,{self.H},{self.W},{self.in_channel}], got {tuple(x.shape)}")
        if len(mean) != self.in_channel or len(std) != self.in_channel:
            raise ValueError("mean / std need one value per input channel")
        x = x.contiguous()
        B = x.shape[0]
        if out is None:
            out = torch.empty((B, 1, self.H, self.W), device=x.device, dtype=torch.float32)
        m = (C.c_float * self.in_channel)(*[float(v) for v in mean])
        s = (C.c_float * self.in_channel)(*[float(v) for v in std])
        with torch.cuda.device(x.device):
            st = torch.cuda.current_stream().cuda_stream
            _lib.check(self._l.isg_forward_u8(self._h, x.data_ptr(), B, m, s, out.data_ptr(), st))
        return out

    # ---- introspection ----
    @property
    def workspace_bytes(self):
        return int(self._l.isg_plan_workspace_bytes(self._h))

    @property
    def launches_per_forward(self):
        return int(self._l.isg_plan_launches_per_forward(self._h))

    @property
    def macs_per_image(self):
        return int(self._l.isg_plan_macs_per_image(self._h))

    def taps(self, n):
        """Intermediate activations of the last forward group as fp32 NCHW CPU tensors (tests)."""
        out = {}
        for i in range(self._l.isg_plan_tap_count(self._h)):
            c, h, w = C.c_int(), C.c_int(), C.c_int()
            _lib.check(self._l.isg_plan_tap_shape(self._h, i, C.byref(c), C.byref(h), C.byref(w)))
            t = torch.empty((n, c.value, h.value, w.value), dtype=torch.float32)
            _lib.check(self._l.isg_plan_tap_read(self._h, i, n, t.data_ptr()))
            out[self._l.isg_plan_tap_name(self._h, i).decode()] = t
        return out
