"""Depthwise convolutions of the bottleneck blocks (model/segment.py:58, :93-96, :121 -- ``nn.Conv2d(groups=channels)``:
3x3 with dilation 1/2/4, (5,1) and (1,5)) for the training step: forward, input gradient and weight/bias gradient in
libisg (``isg_dwconv_forward`` / ``isg_dwconv_wgrad``).  ``DepthwiseConv2d`` is an ``nn.Conv2d`` subclass with the same
``state_dict`` keys; eval mode is lowered by the engine and never comes here."""
import os

import torch
from torch import nn

from . import _lib


def _ptr(t):
    return t.data_ptr() if t is not None else None


class _DwConv(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, padding, dilation):
        x = x.contiguous()
        B, C, H, W = x.shape
        kh, kw = weight.shape[2], weight.shape[3]
        w32 = weight.detach().to(torch.float32).contiguous()
        b32 = bias.detach().to(torch.float32).contiguous() if bias is not None else None
        y = torch.empty_like(x)
        with torch.cuda.device(x.device):
            _lib.check(_lib.lib().isg_dwconv_forward(
                x.data_ptr(), int(x.dtype == torch.bfloat16), B, C, H, W, w32.data_ptr(), _ptr(b32), kh, kw,
                padding[0], padding[1], dilation[0], dilation[1], 0, y.data_ptr(), torch.cuda.current_stream().cuda_stream))
        ctx.save_for_backward(x, w32)
        ctx.geom = (kh, kw, padding, dilation, bias is not None, weight.dtype, weight.shape)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w32 = ctx.saved_tensors
        kh, kw, padding, dilation, has_bias, wdtype, wshape = ctx.geom
        B, C, H, W = x.shape
        dy = dy.contiguous()
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        l, bf = _lib.lib(), int(x.dtype == torch.bfloat16)
        dx = dw = db = None
        with torch.cuda.device(x.device):
            st = torch.cuda.current_stream().cuda_stream
            if ctx.needs_input_grad[0]:
                dx = torch.empty_like(x)
                _lib.check(l.isg_dwconv_forward(dy.data_ptr(), bf, B, C, H, W, w32.data_ptr(), None, kh, kw, padding[0],
                                                padding[1], dilation[0], dilation[1], 1, dx.data_ptr(), st))
            if ctx.needs_input_grad[1] or (has_bias and ctx.needs_input_grad[2]):
                dw = torch.empty(wshape, dtype=torch.float32, device=x.device)
                db = torch.empty(C, dtype=torch.float32, device=x.device) if has_bias else None
                ws = torch.empty(l.isg_dwconv_wgrad_workspace_bytes(C) // 4, dtype=torch.float32, device=x.device)
                _lib.check(l.isg_dwconv_wgrad(dy.data_ptr(), x.data_ptr(), bf, B, C, H, W, kh, kw, padding[0], padding[1],
                                              dilation[0], dilation[1], dw.data_ptr(), _ptr(db), ws.data_ptr(), st))
                if wdtype != torch.float32:
                    dw = dw.to(wdtype)
        return dx, dw, db, None, None


class DepthwiseConv2d(nn.Conv2d):
    """``nn.Conv2d`` with ``groups == in_channels == out_channels``; the train-mode CUDA path (stride 1, same-size output,
    bf16 / fp32) runs in libisg, anything else (eval, CPU, ``ISG_DW_LIBRARY=1``) is the parent class."""

    def native(self, x):
        if os.environ.get("ISG_DW_LIBRARY"):
            return False
        k, p, d = self.kernel_size, self.padding, self.dilation
        return bool(self.training and x.is_cuda and x.dim() == 4 and x.dtype in (torch.float32, torch.bfloat16)
                    and self.groups == self.in_channels == self.out_channels and self.stride == (1, 1)
                    and self.padding_mode == "zeros" and not isinstance(p, str)
                    and 2 * p[0] == d[0] * (k[0] - 1) and 2 * p[1] == d[1] * (k[1] - 1) and k[0] * k[1] <= 25
                    and x.numel() > 0)

    def conv(self, x, with_bias=True):
        bias = self.bias if with_bias else None
        if self.native(x):
            return _DwConv.apply(x, self.weight, bias, self.padding, self.dilation)
        return nn.functional.conv2d(x, self.weight, bias, self.stride, self.padding, self.dilation, self.groups)

    def forward(self, x):
        return self.conv(x, True)


class _ConvTk2s(torch.autograd.Function):
    """forward in libisg (``isg_convt_k2s_forward``); backward = the library's ``convolution_backward`` of the same op"""

    @staticmethod
    def forward(ctx, x, weight, bias, stride, padding):
        x = x.contiguous()
        B, Cin, H, W = x.shape
        Cout, k = weight.shape[1], weight.shape[2]
        w32 = weight.detach().to(torch.float32).contiguous()
        b32 = bias.detach().to(torch.float32).contiguous() if bias is not None else None
        out = torch.empty((B, Cout, (H - 1) * stride - 2 * padding + k, (W - 1) * stride - 2 * padding + k),
                          dtype=x.dtype, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(_lib.lib().isg_convt_k2s_forward(
                x.data_ptr(), int(x.dtype == torch.bfloat16), B, Cin, Cout, H, W, k, stride, padding, w32.data_ptr(),
                _ptr(b32), out.data_ptr(), torch.cuda.current_stream().cuda_stream))
        ctx.save_for_backward(x, weight)
        ctx.geom = (stride, padding, bias is not None)
        return out

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        stride, padding, has_bias = ctx.geom
        w = weight.detach().to(x.dtype)
        dy = dy.contiguous().to(x.dtype)
        mask = [ctx.needs_input_grad[0], ctx.needs_input_grad[1], bool(has_bias and ctx.needs_input_grad[2])]
        dx, dw, db = torch.ops.aten.convolution_backward(
            dy, x, w, [weight.shape[1]] if has_bias else None, [stride, stride], [padding, padding], [1, 1], True, [0, 0], 1, mask)
        if dw is not None:
            dw = dw.to(weight.dtype)
        if db is not None:
            db = db.to(weight.dtype)
        return dx, dw, db, None, None


class ConvTranspose2dK2S(nn.ConvTranspose2d):
    """``nn.ConvTranspose2d`` with ``kernel == 2 * stride`` (segment.py:435 ``bottle6_1``); train-mode CUDA forward in
    libisg when it fits the kernel's domain, otherwise (and ``ISG_CT_LIBRARY=1``) the parent class."""

    def native(self, x):
        if os.environ.get("ISG_CT_LIBRARY"):
            return False
        k, s, p = self.kernel_size, self.stride, self.padding
        return bool(self.training and x.is_cuda and x.dim() == 4 and x.dtype in (torch.float32, torch.bfloat16)
                    and k[0] == k[1] and s[0] == s[1] and p[0] == p[1] and k[0] == 2 * s[0] and 0 <= p[0] < s[0]
                    and self.groups == 1 and self.dilation == (1, 1) and self.output_padding == (0, 0)
                    and self.out_channels <= 8 and self.in_channels * self.out_channels * k[0] * k[0] * 4 <= 40 * 1024
                    and x.numel() > 0)

    def forward(self, x, output_size=None):
        if output_size is None and self.native(x):
            return _ConvTk2s.apply(x, self.weight, self.bias, self.stride[0], self.padding[0])
        return super().forward(x, output_size)
