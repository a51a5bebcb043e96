"""Train-mode BatchNorm2d of the reference's ``Conv`` block (model/segment.py:41; ``model.train()`` at
train_instance.py:372) in libisg: batch statistics, normalisation, running-stat update, the activation that
follows it in the block (segment.py:42-45) and the backward of both, as four streaming kernels
(``isg_bn_train_forward`` / ``isg_bn_train_backward``).  Same module class, same ``state_dict`` keys; eval mode
is folded into the conv weights by the engine and never comes here."""
import os

import torch
import torch.nn.functional as F
from torch import nn

from . import _lib

ACT_NONE, ACT_RELU, ACT_PRELU = 0, 1, 2


def _workspace(x, channels):
    return torch.empty(channels * _lib.lib().isg_bn_max_split() * 3, dtype=torch.float64, device=x.device)


def _ptr(t):
    return t.data_ptr() if t is not None else None


class _BnTrain(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, momentum, eps, mean_offset, act, slope):
        B, C = x.shape[0], x.shape[1]
        hw = x.numel() // (B * C)
        x = x.contiguous()
        y = torch.empty_like(x)
        mean = torch.empty(C, dtype=torch.float32, device=x.device)
        invstd = torch.empty(C, dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(_lib.lib().isg_bn_train_forward(
                x.data_ptr(), int(x.dtype == torch.bfloat16), B, C, hw, _ptr(weight), _ptr(bias), _ptr(mean_offset),
                _ptr(running_mean), _ptr(running_var), float(momentum), float(eps), int(act), _ptr(slope),
                slope.numel() if slope is not None else 0, y.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                _workspace(x, C).data_ptr(), torch.cuda.current_stream().cuda_stream))
        ctx.save_for_backward(x, weight, bias, mean, invstd, slope)
        ctx.act = int(act)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight, bias, mean, invstd, slope = ctx.saved_tensors
        B, C = x.shape[0], x.shape[1]
        hw = x.numel() // (B * C)
        dy = dy.contiguous()
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        need = ctx.needs_input_grad
        dev = x.device
        dx = torch.empty_like(x) if need[0] else None
        dgamma = torch.empty(C, dtype=torch.float32, device=dev) if (need[1] and weight is not None) else None
        dbeta = torch.empty(C, dtype=torch.float32, device=dev) if (need[2] and bias is not None) else None
        dslope = torch.empty(C, dtype=torch.float32, device=dev) if (need[9] and slope is not None) else None
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().isg_bn_train_backward(
                dy.data_ptr(), x.data_ptr(), int(x.dtype == torch.bfloat16), B, C, hw, _ptr(weight), _ptr(bias),
                mean.data_ptr(), invstd.data_ptr(), ctx.act, _ptr(slope), slope.numel() if slope is not None else 0,
                _ptr(dx), _ptr(dgamma), _ptr(dbeta), _ptr(dslope), _workspace(x, C).data_ptr(),
                torch.cuda.current_stream().cuda_stream))
        if dslope is not None and slope.numel() == 1:
            dslope = dslope.sum().reshape(slope.shape)
        return dx, dgamma, dbeta, None, None, None, None, None, None, dslope


_DEFERRED_COUNTERS = None      # list while a deferred_counters() block is active


class deferred_counters:
    """Inside the block every native BatchNorm2d records its ``num_batches_tracked`` instead of launching its own
    ``add_(1)``; leaving the block adds 1 to all of them with one multi-tensor launch (74 launches -> 1 per step)."""

    def __enter__(self):
        global _DEFERRED_COUNTERS
        self.prev, _DEFERRED_COUNTERS = _DEFERRED_COUNTERS, []
        return self

    def __exit__(self, *exc):
        global _DEFERRED_COUNTERS
        todo, _DEFERRED_COUNTERS = _DEFERRED_COUNTERS, self.prev
        if todo and exc[0] is None:
            torch._foreach_add_(todo, 1)
        return False


class BatchNorm2d(nn.BatchNorm2d):
    """``nn.BatchNorm2d`` whose train-mode CUDA path (bf16 / fp32, fixed momentum) runs in libisg.  Everything else
    -- eval mode, CPU tensors (the boundary tests compare train mode against the reference on CPU), cumulative
    averaging (``momentum=None``), other dtypes -- is the parent class unchanged."""

    def native(self, x):
        if os.environ.get("ISG_BN_LIBRARY"):      # A/B switch: the parent class's library kernels
            return False
        return bool(self.training and x.is_cuda and x.dim() == 4 and self.momentum is not None
                    and self.track_running_stats and x.dtype in (torch.float32, torch.bfloat16) and x.numel() > 0)

    def forward(self, x):
        if self.native(x):
            return self.fused(x)
        return super().forward(x)

    def fused(self, x, mean_offset=None, act=None):
        """``act(bn(x + mean_offset))`` for a native ``x``: ``mean_offset`` is a per-channel constant left out of ``x``
        (the conv bias: train-mode normalisation cancels it, only ``running_mean`` sees it; its gradient is exactly
        zero) and ``act`` the block's ``nn.PReLU`` / ``nn.ReLU`` / ``nn.Identity`` module, fused into the same kernels."""
        kind, slope = ACT_NONE, None
        if isinstance(act, nn.PReLU):
            kind, slope = ACT_PRELU, act.weight
        elif isinstance(act, nn.ReLU):
            kind = ACT_RELU
        elif act is not None and not isinstance(act, nn.Identity):
            raise TypeError(f"cannot fuse activation {type(act).__name__}")
        if _DEFERRED_COUNTERS is not None:
            _DEFERRED_COUNTERS.append(self.num_batches_tracked)
        else:
            self.num_batches_tracked.add_(1)
        off = mean_offset.detach() if mean_offset is not None else None
        return _BnTrain.apply(x, self.weight, self.bias, self.running_mean, self.running_var, self.momentum, self.eps,
                              off, kind, slope)


def conv_bn_act(conv, bn, act, x):
    """The reference's ``Conv.forward`` (segment.py:47-48): ``act(bn(conv(x)))``.  On the native train path the conv
    runs without its bias (it cancels in the normalisation; ``ISG_TRAIN_KEEP_CONV_BIAS=1`` keeps the literal
    sequence) and BatchNorm + activation are one forward and one backward kernel pair."""
    if (isinstance(bn, BatchNorm2d) and bn.training and x.is_cuda and bn.momentum is not None and bn.track_running_stats
            and conv.padding_mode == "zeros" and not os.environ.get("ISG_TRAIN_KEEP_CONV_BIAS")):
        if hasattr(conv, "conv"):            # dwconv.DepthwiseConv2d: native stencil kernels
            z = conv.conv(x, with_bias=False)
        else:
            z = F.conv2d(x, conv.weight, None, conv.stride, conv.padding, conv.dilation, conv.groups)
        if bn.native(z):
            return bn.fused(z, conv.bias, act)
        if conv.bias is not None:
            z = z + conv.bias.view(1, -1, 1, 1).to(z.dtype)
        return act(bn(z))
    return act(bn(conv(x)))
