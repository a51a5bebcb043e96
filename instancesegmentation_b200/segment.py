"""Drop-in ``Segment`` network (reference: model/segment.py:347-534).

The module tree below exists to carry parameters: it reproduces the reference's
588 ``state_dict`` keys (names, shapes, dtypes, order) and its seeded
initialisation bit for bit, so checkpoints written by the reference
(train_instance.py:497-507) load with ``strict=True``.

Inference (``eval()`` mode, CUDA tensor, no autograd) never touches the torch
layers: the whole forward runs in ``libisg.so`` (hand-written sm_100a kernels,
``include/isg.h``).  There is no CPU implementation of that path and no
fallback: a missing library or a CPU tensor in eval mode raises.

Training mode (``train()``; reference train_instance.py:372-380) needs autograd
and batch-statistics BatchNorm, so it evaluates the same parameter tree through
torch's differentiable ops -- SURVEY.md section 7.3-E, "first cut"; gradients
are exchanged by ``instancesegmentation_b200.distributed``.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .bn import BatchNorm2d, conv_bn_act
from .dwconv import ConvTranspose2dK2S, DepthwiseConv2d

__all__ = ["Segment", "topology"]


# --------------------------------------------------------------------------
# topology table (one row per top-level attribute of the reference network)
# --------------------------------------------------------------------------
def topology(in_channel: int):
    """Rows ``(attribute, kind, params)`` in registration order (segment.py:357-438)."""
    stem_out = 16 + in_channel
    res3 = lambda c, m, d=1: ("res3", dict(c=c, mid=m, dil=d))
    res5 = lambda c, m: ("res5", dict(c=c, mid=m))
    wide = [res3(128, 48), res3(128, 48, 2), res3(128, 48), res3(128, 48, 4), res5(128, 48)]
    return [
        ("init_conv", "stem", dict(cin=in_channel, mid=16, cout=stem_out)),
        ("bottle1_1", "down", dict(cin=stem_out, mid=16, cout=48)),
        ("bottle1_x", "seq", [res3(48, 16)] * 4),
        ("bottle2_1", "down", dict(cin=48, mid=16, cout=128)),
        ("bottle2_x", "seq", wide),
        ("bottle3_1", "proj", dict(cin=256, mid=48, cout=128, last="prelu")),
        ("bottle3_x", "seq", wide),
        ("bottle4_1up", "up", dict(cin=128, mid=16, cout=48, skip=48)),
        ("bottle4_2", "proj", dict(cin=96, mid=16, cout=48, last="relu")),
        ("bottle4_3", "plain", dict(c=48, mid=16)),
        ("bottle5_1up", "up", dict(cin=48, mid=4, cout=16, skip=stem_out, redraw=True)),
        ("bottle5_2", "plain", dict(c=16, mid=4)),
        ("bottle6_1", "head_up", dict(cin=16, cout=4)),
        ("bottle6_2", "head_out", dict(cin=4, cout=1)),
    ]


def _act(kind, ch):
    if kind == "prelu":
        return nn.PReLU(ch)
    if kind == "relu":
        return nn.ReLU(inplace=True)
    return nn.Identity()


class _Unit(nn.Module):
    """conv(+bias) -> BatchNorm -> activation; keys ``conv.* bn.* act.*`` (segment.py:34-45)."""

    def __init__(self, cin, cout, k=1, s=1, p=None, g=1, d=1, act=None):
        super().__init__()
        if p is None:
            p = k // 2 if isinstance(k, int) else tuple(v // 2 for v in k)
        a = _act(act, cout)  # constructed first, like the reference's call sites
        conv_cls = DepthwiseConv2d if (g > 1 and g == cin == cout) else nn.Conv2d
        self.conv = conv_cls(cin, cout, k, s, p, groups=g, dilation=d, bias=True)
        self.bn = BatchNorm2d(cout)
        self.act = a

    def forward(self, x):
        return conv_bn_act(self.conv, self.bn, self.act, x)


def _seq(*mods):
    return nn.Sequential(*mods)


class _Block(nn.Module):
    """One residual block of the reference; ``kind`` selects wiring and child names."""

    def __init__(self, kind, **a):
        super().__init__()
        self.kind = kind
        if kind == "stem":  # init_head_s4, segment.py:19-31
            self.layer1 = _Unit(a["cin"], a["mid"], 5, 2, 2, act="prelu")
            self.layer2 = _Unit(a["mid"], a["cout"] - a["cin"], 5, 2, 2, act="prelu")
        elif kind == "res3":  # Bottleneck3x3, segment.py:52-79
            c, m, d = a["c"], a["mid"], a["dil"]
            self.convs = _seq(_Unit(c, m, 1, act="prelu"),
                              _Unit(m, m, 3, p=d, d=d, g=m, act="prelu"),
                              _Unit(m, c, 1))
            self.prelu = nn.PReLU(c)
        elif kind == "res5":  # Bottleneck5x5, segment.py:82-111
            c, m = a["c"], a["mid"]
            self.convs = _seq(_Unit(c, m, 1, act="prelu"),
                              DepthwiseConv2d(m, m, (5, 1), padding=(2, 0), groups=m),
                              _Unit(m, m, (1, 5), p=(0, 2), g=m, act="prelu"),
                              _Unit(m, c, 1))
            self.prelu = nn.PReLU(c)
        elif kind == "down":  # BottleneckDown2, segment.py:114-150
            ci, m, co = a["cin"], a["mid"], a["cout"]
            self.convs = _seq(_Unit(ci, m, 2, 2, 0, act="prelu"),
                              _Unit(m, m, 3, 1, 1, g=m, act="prelu"),
                              _Unit(m, co, 1))
            self.convm = _seq(_Unit(ci, co, 1))
            self.prelu = nn.PReLU(co)
        elif kind == "proj":  # BottleneckDim_Res, segment.py:153-209 (PReLU inside both variants)
            ci, m, co = a["cin"], a["mid"], a["cout"]
            self.last = a["last"]
            self.convs = _seq(_Unit(ci, m, 1, act="prelu"),
                              _Unit(m, m, 3, p=1, g=m, act="prelu"),
                              _Unit(m, co, 1))
            self.resconv = _seq(_Unit(ci, co, 1))
            self.prelu = nn.PReLU(co)  # unused (but present) when last == "relu"
            self.relu = nn.ReLU(inplace=True)
        elif kind == "plain":  # BottleneckDim(usePrelu=False), segment.py:232-261
            c, m = a["c"], a["mid"]
            self.convs = _seq(_Unit(c, m, 1, act="relu"),
                              _Unit(m, m, 3, p=1, act="relu"),
                              _Unit(m, c, 1))
            self.prelu = nn.PReLU(c)  # never used by forward; kept for strict loading
            self.relu = nn.ReLU(inplace=True)
        elif kind == "up":  # BottleneckUp_Res(_Other), segment.py:296-344
            ci, m, co, sk = a["cin"], a["mid"], a["cout"], a["skip"]
            self.convs = _seq(_Unit(ci, m, 1, act="relu"),
                              nn.ConvTranspose2d(m, m, 4, stride=2, padding=1),
                              BatchNorm2d(m),
                              nn.ReLU(inplace=True),
                              _Unit(m, co, 1))
            self.conv2 = _seq(_Unit(ci, co, 1))
            mix = nn.Conv2d(2 * co, co, 1)
            if a.get("redraw"):
                # the reference subclass builds the base 1x1 first and then replaces it
                # (segment.py:340-344); repeating that keeps the seeded RNG stream aligned
                mix = nn.Conv2d(co + sk, co, 1)
            self.uppool = _seq(nn.UpsamplingNearest2d(scale_factor=2), mix)
        else:
            raise ValueError(kind)

    # ---- differentiable torch path (training only) ----
    def forward(self, x, skip=None):
        k = self.kind
        if k == "stem":
            return torch.cat((F.max_pool2d(x, 4, 4), self.layer2(self.layer1(x))), 1)
        if k in ("res3", "res5"):
            return self.prelu(self.convs(x) + x)
        if k == "down":
            pooled = F.max_pool2d(x, 2, 2)
            return self.prelu(self.convs(x) + self.convm(pooled)), pooled
        if k == "proj":
            y = self.convs(x) + self.resconv(x)
            return self.prelu(y) if self.last == "prelu" else self.relu(y)
        if k == "plain":
            return self.relu(self.convs(x) + x)
        if k == "up":
            side = self.uppool(torch.cat((self.conv2(x), skip), 1))
            return F.relu(self.convs(x) + side)
        raise AssertionError(k)


def _make(kind, p):
    if kind == "seq":
        return nn.Sequential(*[_Block(k, **a) for k, a in p])
    if kind == "head_up":
        return ConvTranspose2dK2S(p["cin"], p["cout"], kernel_size=8, padding=2, stride=4)
    if kind == "head_out":
        return nn.Conv2d(p["cin"], p["cout"], kernel_size=3, padding=1)
    return _Block(kind, **p)


class Segment(nn.Module):
    """``Segment(in_channel)``; ``forward(x[B,in_channel,H,W]) -> logits[B,1,H,W]``.

    Extra (non-reference) attributes: ``precision`` ("fp32" | "bf16", kernel
    arithmetic of the CUDA path; default "fp32" = reference numerics within 1e-4)
    and ``max_batch`` (engine workspace is sized for this many images per launch
    group; larger batches are processed in groups).
    """

    def __init__(self, in_channel):
        super().__init__()
        self.export = False
        self.output_mid_features = False
        self.in_channel = int(in_channel)
        self.init_Dim = 16 + in_channel
        self.bottle1_downDim, self.bottle1_Dim = 16, 48
        self.bottle2_downDim, self.bottle2_Dim = 48, 128
        for name, kind, p in topology(in_channel):
            setattr(self, name, _make(kind, p))
        self.weights_init()
        self.precision = "fp32"
        self.max_batch = 64
        self._engines = {}
        self._weights_epoch = 0

    def weights_init(self):
        """Kaiming-normal on every Conv2d, zero bias, BN (1, 0) -- segment.py:451-464."""
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_in", nonlinearity="relu")
                if m.bias is not None:
                    nn.init.zeros_(m.bias)
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.ones_(m.weight)
                nn.init.zeros_(m.bias)

    # ---- bookkeeping: anything that may change parameters drops the packed copy ----
    def invalidate(self):
        """Forget packed device weights (call after editing parameters in place)."""
        self._weights_epoch += 1

    def train(self, mode=True):
        self.invalidate()
        return super().train(mode)

    def load_state_dict(self, *a, **k):
        self.invalidate()
        return super().load_state_dict(*a, **k)

    def _apply(self, fn, *a, **k):
        self.invalidate()
        return super()._apply(fn, *a, **k)

    def __getstate__(self):
        st = dict(self.__dict__)
        st["_engines"] = {}
        return st

    def __deepcopy__(self, memo):
        import copy
        new = self.__class__.__new__(self.__class__)
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            new.__dict__[k] = {} if k == "_engines" else copy.deepcopy(v, memo)
        return new

    @staticmethod
    def check_size(h, w):
        """The reference works iff floor(H/4) % 4 == 0; we accept the
        sizes where its output equals its input size, i.e. multiples of 16."""
        if h < 16 or w < 16 or h % 16 or w % 16:
            raise ValueError(f"Segment: H and W must be multiples of 16 (>=16), got {h}x{w}")

    # ---- forward ----
    def forward(self, x):
        if x.dim() != 4 or x.shape[1] != self.in_channel:
            raise ValueError(f"Segment: expected [B,{self.in_channel},H,W], got {tuple(x.shape)}")
        self.check_size(x.shape[2], x.shape[3])
        if self.training or (torch.is_grad_enabled() and x.requires_grad):
            return self._forward_autograd(x)
        return self._forward_engine(x)

    def _forward_autograd(self, x):
        s = self.init_conv(x)
        d1, keep1 = self.bottle1_1(s)
        d2, keep2 = self.bottle2_1(self.bottle1_x(d1))
        t = self.bottle3_x(self.bottle3_1(torch.cat((self.bottle2_x(d2), d2), 1)))
        t = self.bottle4_1up(t, keep2)
        t = self.bottle4_3(self.bottle4_2(torch.cat((d1, t), 1)))
        t = self.bottle5_2(self.bottle5_1up(t, keep1))
        return self.bottle6_2(self.bottle6_1(t))

    def forward_u8(self, x, mean=(0.5, 0.5, 0.5), std=(0.5, 0.5, 0.5), out=None):
        """Eval-mode forward from uint8 NHWC images [B,H,W,C] on the device: the reference's
        ``ToTensor`` + ``Normalize(mean, std)`` (train_instance.py:80-85) run inside the engine's first
        kernel.  Equals ``forward(((x / 255 - mean) / std).permute(0, 3, 1, 2))``."""
        if x.dim() != 4 or x.shape[3] != self.in_channel:
            raise ValueError(f"Segment: expected uint8 [B,H,W,{self.in_channel}], got {tuple(x.shape)}")
        self.check_size(x.shape[1], x.shape[2])
        if not x.is_cuda:
            raise RuntimeError("Segment.forward_u8 runs only in the sm_100a CUDA engine (libisg.so)")
        eng = self._engine_for(x.device, x.shape[0], x.shape[1], x.shape[2])
        return eng.forward_u8(x, mean, std, out=out)

    def _forward_engine(self, x):
        if not x.is_cuda:
            raise RuntimeError(
                "Segment: eval-mode forward runs only in the sm_100a CUDA engine (libisg.so); "
                "got a CPU tensor and there is no CPU fallback")
        if x.dtype != torch.float32:
            raise TypeError(f"Segment: input must be float32, got {x.dtype}")
        B, _, H, W = x.shape
        return self._engine_for(x.device, B, H, W).forward(x)

    def _engine_for(self, device, B, H, W):
        from .engine import SegmentEngine
        # one plan per (device, size, precision): a plan serves any batch up to its max_batch (isg_forward groups
        # larger ones), so a new one -- autotune + its own workspace -- is only built when B exceeds the cached
        # plan's capacity (and that one is then dropped), never per distinct batch size
        key = (device.index, H, W, self.precision)
        need = min(self.max_batch, max(B, 1))
        eng = self._engines.get(key)
        if eng is None or eng.max_batch < need:
            eng = SegmentEngine(device, self.in_channel, H, W, need, self.precision)
            self._engines[key] = eng
        if eng.weights_epoch != self._weights_epoch:
            eng.load_state_dict(self.state_dict())
            eng.weights_epoch = self._weights_epoch
        return eng

    def train_batch(self, x, heatmaps=None):
        """sigmoid(forward(cat([x, heatmaps], 1))) -- segment.py:531-534; ``heatmaps``
        may be omitted (train_instance.py:377 calls it that way)."""
        if heatmaps is not None:
            x = torch.cat([x, heatmaps], dim=1)
        return torch.sigmoid(self.forward(x))
