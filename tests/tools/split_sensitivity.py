"""TEST TOOLING (CPU): which operand precision does each conv of Segment need so that the logits stay within
the north_star tolerance (max-abs 2e-2 / mean-abs 2e-3)?

Emulates the tensor-core engine's arithmetic on the CPU on top of the oracle's graph: every conv runs in
float64 on operands that were first rounded the way the engine rounds them --
  "bf16"  one bf16 value per operand element (the plain ISG_MODE_BF16 path),
  "x3"    split precision: a = a_hi + a_lo, w = w_hi + w_lo (two bf16 values each); the product keeps
          a_hi w_hi + a_lo w_hi + a_hi w_lo (three MMAs, the lo*lo term is dropped),
  "x2a"   only the activation is split (a_hi w_hi + a_lo w_hi), "x2w" only the weight,
  "fp32"  no rounding --
and every activation handed from one kernel to the next is stored as bf16 ("bf16") or as a hi+lo pair.
Prints the logit error per configuration.  Usage: python tests/tools/split_sensitivity.py [HxW]
"""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from instancesegmentation_b200 import synth  # noqa: E402
from oracle.segment_oracle import segment_forward  # noqa: E402

EPS = 1e-5


def bf(x):
    return x.float().bfloat16().double()


def split(x):
    hi = bf(x)
    return hi, bf(x.double() - hi)


def f16(x):
    return x.float().half().double()


def qstore(x, mode):
    if mode == "bf16":
        return bf(x)
    if mode == "fp16":
        return f16(x)
    if mode == "fp32":
        return x.float().double()
    hi, lo = split(x)
    return hi + lo


class Emu:
    def __init__(self, sd, layer_mode, store_mode, dw_store=None):
        self.sd = {k: v.double() if v.dtype.is_floating_point else v for k, v in sd.items()}
        self.layer_mode = layer_mode      # fn(name) -> operand mode of that conv
        self.store_mode = store_mode      # fn(name) -> storage mode of that unit's output
        self.dw_store = dw_store or store_mode

    def fold(self, p, bn=None, conv=None):
        sd = self.sd
        conv, bn = conv or p + ".conv", bn or p + ".bn"
        s = sd[bn + ".weight"] / torch.sqrt(sd[bn + ".running_var"] + EPS)
        return s, sd[bn + ".bias"] + (sd[conv + ".bias"] - sd[bn + ".running_mean"]) * s

    def mm(self, name, x, w, fn):
        """fn(x, w) = the linear conv; operands rounded per the layer's mode"""
        mode = self.layer_mode(name)
        if mode == "fp32":
            return fn(x, w)
        if mode == "bf16":
            return fn(bf(x), bf(w))
        if mode == "fp16":        # one pass with fp16 operands (11 significand bits)
            return fn(f16(x), f16(w))
        if mode == "fp16x2":      # fp16 activation pair hi + lo, fp16 weight: A_hi.W + A_lo.W
            xh = f16(x)
            return fn(xh, f16(w)) + fn(f16(x.double() - xh), f16(w))
        xh, xl = split(x)
        wh, wl = split(w)
        y = fn(xh, wh)
        if mode in ("x3", "x2a"):
            y = y + fn(xl, wh)
        if mode in ("x3", "x2w"):
            y = y + fn(xh, wl)
        return y

    def unit(self, p, x, act, dw=False, **kw):
        sd = self.sd
        s, sh = self.fold(p)
        w = sd[p + ".conv.weight"] * s.view(-1, 1, 1, 1)
        if dw:   # depthwise stencils run in fp32 on CUDA cores on the stored operand
            y = F.conv2d(x, w, None, **kw).float().double()
        else:
            y = self.mm(p, x, w, lambda a, b: F.conv2d(a, b, None, **kw)).float().double()
        y = y + sh.view(1, -1, 1, 1)
        if act == "prelu":
            return F.prelu(y, sd[p + ".act.weight"])
        return F.relu(y) if act == "relu" else y

    def st(self, name, x):
        return qstore(x, self.store_mode(name))

    def dwpart(self, p, y, d=None, five=False):
        """first conv output -> depthwise part of a bottleneck; intermediates stored per dw_store"""
        sd = self.sd
        y = qstore(y, self.dw_store(p + ".m1"))
        if five:
            m = sd[p + ".convs.1.weight"].shape[0]
            y = F.conv2d(y, sd[p + ".convs.1.weight"], sd[p + ".convs.1.bias"], padding=(2, 0), groups=m)
            y = self.unit(p + ".convs.2", y, "prelu", dw=True, padding=(0, 2), groups=m)
        else:
            m = sd[p + ".convs.1.conv.weight"].shape[0]
            y = self.unit(p + ".convs.1", y, "prelu", dw=True, padding=d, dilation=d, groups=m)
        return qstore(y, self.dw_store(p + ".d"))

    def res(self, p, x, d, five=False):
        y = self.dwpart(p, self.unit(p + ".convs.0", x, "prelu"), d, five)
        y = self.unit(p + (".convs.3" if five else ".convs.2"), y, None)
        return self.st(p, F.prelu(y + x, self.sd[p + ".prelu.weight"]))

    def down(self, p, x):
        y = self.dwpart(p, self.unit(p + ".convs.0", x, "prelu", stride=2), 1)
        y = self.unit(p + ".convs.2", y, None)
        pooled = F.max_pool2d(x, 2, 2)
        return self.st(p, F.prelu(y + self.unit(p + ".convm.0", pooled, None), self.sd[p + ".prelu.weight"])), pooled

    def proj(self, p, x, last):
        y = self.dwpart(p, self.unit(p + ".convs.0", x, "prelu"), 1)
        y = self.unit(p + ".convs.2", y, None) + self.unit(p + ".resconv.0", x, None)
        return self.st(p, F.prelu(y, self.sd[p + ".prelu.weight"]) if last == "prelu" else F.relu(y))

    def plain(self, p, x):
        y = self.st(p + ".m1", self.unit(p + ".convs.0", x, "relu"))
        y = self.st(p + ".m2", self.unit(p + ".convs.1", y, "relu", padding=1))
        return self.st(p, F.relu(self.unit(p + ".convs.2", y, None) + x))

    def up(self, p, x, skip):
        sd = self.sd
        y = self.st(p + ".m1", self.unit(p + ".convs.0", x, "relu"))
        s, sh = self.fold(p, bn=p + ".convs.2", conv=p + ".convs.1")
        w = sd[p + ".convs.1.weight"] * s.view(1, -1, 1, 1)
        y = self.mm(p + ".convs.1", y, w, lambda a, b: F.conv_transpose2d(a, b, None, stride=2, padding=1)).float().double()
        y = self.st(p + ".m2", F.relu(y + sh.view(1, -1, 1, 1)))
        y = self.unit(p + ".convs.4", y, None)
        side = torch.cat((self.st(p + ".c2", self.unit(p + ".conv2.0", x, None)), skip), 1)
        side = self.mm(p + ".uppool.1", side, sd[p + ".uppool.1.weight"], lambda a, b: F.conv2d(a, b, None)).float().double()
        side = self.st(p + ".side", side + sd[p + ".uppool.1.bias"].view(1, -1, 1, 1))
        return self.st(p, F.relu(y + F.interpolate(side, scale_factor=2, mode="nearest")))

    def wide(self, p, x):
        for i, d in enumerate((1, 2, 1, 4)):
            x = self.res(f"{p}.{i}", x, d)
        return self.res(f"{p}.4", x, None, five=True)

    @torch.no_grad()
    def forward(self, x):
        sd = self.sd
        x = self.st("input", x.double())
        y = self.st("stem1", self.unit("init_conv.layer1", x, "prelu", stride=2, padding=2))
        y = self.st("stem2", self.unit("init_conv.layer2", y, "prelu", stride=2, padding=2))
        d1, keep1 = self.down("bottle1_1", torch.cat((F.max_pool2d(x, 4, 4), y), 1))
        y = d1
        for i in range(4):
            y = self.res(f"bottle1_x.{i}", y, 1)
        d2, keep2 = self.down("bottle2_1", y)
        y = self.wide("bottle2_x", d2)
        y = self.proj("bottle3_1", torch.cat((y, d2), 1), "prelu")
        y = self.wide("bottle3_x", y)
        y = self.up("bottle4_1up", y, keep2)
        y = self.proj("bottle4_2", torch.cat((d1, y), 1), "relu")
        y = self.plain("bottle4_3", y)
        y = self.up("bottle5_1up", y, keep1)
        y = self.plain("bottle5_2", y)

        def tail(a, w):   # bottle6_1 o bottle6_2 is one linear map of y (no activation in between)
            return F.conv2d(F.conv_transpose2d(a, w, None, stride=4, padding=2), sd["bottle6_2.weight"], None, padding=1)
        out = self.mm("tail", y, sd["bottle6_1.weight"], tail)
        b = F.conv2d(F.conv_transpose2d(torch.zeros_like(y), sd["bottle6_1.weight"], sd["bottle6_1.bias"], stride=4, padding=2),
                     sd["bottle6_2.weight"], sd["bottle6_2.bias"], padding=1)
        return (out + b).float()


def main():
    H, W = (int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "128x128").split("x"))
    z = np.load(os.path.join(ROOT, "tests", "golden", "segment3_state.npz"))
    sd = {k: torch.from_numpy(z[k]) for k in z.files}
    x = torch.from_numpy(synth.batch(2, H, W, seed=77))
    want = segment_forward(sd, x)

    def report(label, lm, sm, dws=None):
        d = (Emu(sd, lm, sm, dws).forward(x) - want).abs()
        print(f"{label:70s} max {float(d.max()):.3e} mean {float(d.mean()):.3e}", flush=True)

    def every(m):
        return lambda name: m
    report("all fp32 (emulation floor)", every("fp32"), every("fp32"))
    report("all bf16 operands, bf16 storage", every("bf16"), every("bf16"))
    report("all x3, hi+lo storage", every("x3"), every("x3"))
    report("all x3, hi+lo storage, dw intermediates (m1, d) bf16", every("x3"), every("x3"), every("bf16"))
    report("all x2a (activation split only), hi+lo storage", every("x2a"), every("x3"))
    report("all x2w (weight split only), hi+lo storage", every("x2w"), every("x3"))
    report("all x3 operands but bf16 storage", every("x3"), every("bf16"))
    report("all fp16 operands, fp16 storage (one MMA pass)", every("fp16"), every("fp16"))
    report("all fp16 operands, hi+lo (bf16 pair) storage", every("fp16"), every("x3"))
    report("fp16x2 (activation fp16 pair, fp16 weights), pair storage", every("fp16x2"), every("x3"))
    if len(sys.argv) > 2 and sys.argv[2] == "fp16only":
        return
    groups = {"stem": "init_conv", "enc8 (bottle1)": "bottle1", "enc16 (bottle2)": "bottle2", "mid16 (bottle3)": "bottle3",
              "dec8 (bottle4)": "bottle4", "dec4 (bottle5)": "bottle5", "tail": "tail"}
    for alt in ("bf16", "x2a", "x2w"):
        for g, pre in groups.items():
            report(f"x3 everywhere except {g}: {alt}", lambda name, pre=pre: alt if name.startswith(pre) else "x3", every("x3"))


if __name__ == "__main__":
    main()
