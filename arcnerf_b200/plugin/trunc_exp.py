# -*- coding: utf-8 -*-
"""Drop-in for arcnerf/ops/trunc_exp.py: exp forward, gradient through exp(clamp(x, -clip, clip)) (trunc_exp.py:7-60),
as two libarcb200 kernels instead of eager torch ops."""
import torch
import torch.nn as nn

from .. import _lib


class TruncExpOps(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, clip):
        xc = x.contiguous()
        if xc.dtype not in (torch.float16, torch.float32):
            xc = xc.float()
        ctx.save_for_backward(xc)
        ctx.clip = clip
        y = torch.empty(xc.shape, dtype=torch.float32, device=xc.device)
        _lib.call('arcb200_trunc_exp_fwd', xc.data_ptr(), int(xc.dtype == torch.float16), 1, y.data_ptr(), xc.numel(), None,
                  _lib.current_stream())
        return y.to(x.dtype)

    @staticmethod
    def backward(ctx, grad):
        (xc,) = ctx.saved_tensors
        g = grad.contiguous().float()
        gx = torch.empty(xc.shape, dtype=torch.float32, device=xc.device)
        _lib.call('arcb200_trunc_exp_bwd', xc.data_ptr(), int(xc.dtype == torch.float16), 1, g.data_ptr(), float(ctx.clip),
                  gx.data_ptr(), xc.numel(), None, _lib.current_stream())
        return gx.to(grad.dtype), None


class TruncExp(nn.Module):
    def __init__(self, clip=15.0):
        super(TruncExp, self).__init__()
        self.clip = clip

    def forward(self, x):
        return TruncExpOps.apply(x, self.clip)
