"""Train-mode BatchNorm2d of the reference's ``Conv`` block (model/segment.py:41; ``model.train()`` at
train_instance.py:372) in libisg: batch statistics, normalisation, running-stat update and the backward as
four streaming kernels (``isg_bn_train_forward`` / ``isg_bn_train_backward``).  Same module class, same
``state_dict`` keys; eval mode is folded into the conv weights by the engine and never comes here."""
import torch
from torch import nn

from . import _lib


def _workspace(x, channels):
    return torch.empty(channels * _lib.lib().isg_bn_max_split() * 2, dtype=torch.float64, device=x.device)


class _BnTrain(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, momentum, eps):
        B, C = x.shape[0], x.shape[1]
        hw = x.numel() // (B * C)
        x = x.contiguous()
        y = torch.empty_like(x)
        mean = torch.empty(C, dtype=torch.float32, device=x.device)
        invstd = torch.empty(C, dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(_lib.lib().isg_bn_train_forward(
                x.data_ptr(), int(x.dtype == torch.bfloat16), B, C, hw,
                weight.data_ptr() if weight is not None else None, bias.data_ptr() if bias is not None else None,
                running_mean.data_ptr() if running_mean is not None else None,
                running_var.data_ptr() if running_var is not None else None,
                float(momentum), float(eps), y.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                _workspace(x, C).data_ptr(), torch.cuda.current_stream().cuda_stream))
        ctx.save_for_backward(x, weight, mean, invstd)
        ctx.has_bias = bias is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight, mean, invstd = ctx.saved_tensors
        B, C = x.shape[0], x.shape[1]
        hw = x.numel() // (B * C)
        dy = dy.contiguous()
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        need_x, need_w, need_b = ctx.needs_input_grad[0], ctx.needs_input_grad[1], ctx.needs_input_grad[2]
        dx = torch.empty_like(x) if need_x else None
        dgamma = torch.empty(C, dtype=torch.float32, device=x.device) if (need_w and weight is not None) else None
        dbeta = torch.empty(C, dtype=torch.float32, device=x.device) if (need_b and ctx.has_bias) else None
        with torch.cuda.device(x.device):
            _lib.check(_lib.lib().isg_bn_train_backward(
                dy.data_ptr(), x.data_ptr(), int(x.dtype == torch.bfloat16), B, C, hw,
                weight.data_ptr() if weight is not None else None, mean.data_ptr(), invstd.data_ptr(),
                dx.data_ptr() if dx is not None else None, dgamma.data_ptr() if dgamma is not None else None,
                dbeta.data_ptr() if dbeta is not None else None, _workspace(x, C).data_ptr(),
                torch.cuda.current_stream().cuda_stream))
        return dx, dgamma, dbeta, None, None, None, None


class BatchNorm2d(nn.BatchNorm2d):
    """``nn.BatchNorm2d`` whose train-mode CUDA path (bf16 / fp32, fixed momentum) runs in libisg.  Everything else
    -- eval mode, CPU tensors (the boundary tests compare train mode against the reference on CPU), cumulative
    averaging (``momentum=None``), other dtypes -- is the parent class unchanged."""

    def forward(self, x):
        if (self.training and x.is_cuda and x.dim() == 4 and self.momentum is not None and self.track_running_stats
                and x.dtype in (torch.float32, torch.bfloat16) and x.numel() > 0):
            self.num_batches_tracked.add_(1)
            return _BnTrain.apply(x, self.weight, self.bias, self.running_mean, self.running_var, self.momentum, self.eps)
        return super().forward(x)
