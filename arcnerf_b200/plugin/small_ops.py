# -*- coding: utf-8 -*-
"""Differentiable building blocks on libarcb200's fp32 kernels (csrc/small_gemm.cu) for the narrow dense layers of the SDF
models: `linear` stands where the reference calls torch.nn.functional.linear through nn.Linear / DenseLayer
(arcnerf/models/base_modules/linear.py), `softplus` where it applies nn.Softplus(beta) (base_modules/activation.py).

Each product is an autograd.Function whose backward is written with the same functions, so autograd can differentiate the
backward again: BaseGeoNet.forward_with_grad (base_network.py:30-44) takes d sdf / d x with create_graph=True and the eikonal
loss then needs the derivative of that gradient w.r.t. the weights.  No cuBLAS: every product is arcb200_gemm_nt / _tn."""
import torch

from .. import _lib

_MAX = 128


def _check(t):
    if not t.is_cuda:
        raise RuntimeError('small_ops need CUDA tensors (no CPU fallback)')
    return t.contiguous().float()


class _MatmulNT(torch.autograd.Function):
    """C (n,N) = A (n,K) @ B (N,K)^T"""

    @staticmethod
    def forward(ctx, a, b):
        a, b = _check(a), _check(b)
        ctx.save_for_backward(a, b)
        n, k = a.shape
        c = torch.empty((n, b.shape[0]), dtype=torch.float32, device=a.device)
        _lib.call('arcb200_gemm_nt', a.data_ptr(), a.stride(0), b.data_ptr(), b.stride(0), None, c.data_ptr(), c.stride(0), n,
                  b.shape[0], k, _lib.current_stream())
        return c

    @staticmethod
    def backward(ctx, dc):
        a, b = ctx.saved_tensors
        da = db = None
        if ctx.needs_input_grad[0]:
            da = _MatmulNT.apply(dc, b.t().contiguous())          # dC @ B
        if ctx.needs_input_grad[1]:
            db = _MatmulTN.apply(dc, a)                           # dC^T @ A
        return da, db


class _MatmulTN(torch.autograd.Function):
    """C (N,K) = A (n,N)^T @ B (n,K): the contraction runs over the long dimension"""

    @staticmethod
    def forward(ctx, a, b):
        a, b = _check(a), _check(b)
        ctx.save_for_backward(a, b)
        c = torch.zeros((a.shape[1], b.shape[1]), dtype=torch.float32, device=a.device)
        _lib.call('arcb200_gemm_tn', a.data_ptr(), a.stride(0), b.data_ptr(), b.stride(0), c.data_ptr(), c.stride(0), a.shape[0],
                  a.shape[1], b.shape[1], _lib.current_stream())
        return c

    @staticmethod
    def backward(ctx, dc):
        a, b = ctx.saved_tensors
        da = db = None
        if ctx.needs_input_grad[0]:
            da = _MatmulNT.apply(b, dc)                           # B @ dC^T   (n,K) x (N,K)^T
        if ctx.needs_input_grad[1]:
            db = _MatmulNT.apply(a, dc.t().contiguous())          # A @ dC     (n,N) x (K,N)^T
        return da, db


def matmul_nt(a, b):
    """a (n,K) @ b (N,K)^T for any N, K: blocks of at most 128 x 128 go to the kernel, partial products are summed"""
    n_out, k_in = b.shape
    if n_out <= _MAX and k_in <= _MAX:
        return _MatmulNT.apply(a, b)
    cols = []
    for j in range(0, n_out, _MAX):
        acc = None
        for k in range(0, k_in, _MAX):
            part = _MatmulNT.apply(a[:, k:k + _MAX], b[j:j + _MAX, k:k + _MAX])
            acc = part if acc is None else acc + part
        cols.append(acc)
    return torch.cat(cols, 1)


def linear(x, weight, bias=None):
    """torch.nn.functional.linear(x, weight, bias) for 2-D x on the library's kernels, differentiable to second order"""
    y = matmul_nt(x, weight)
    return y if bias is None else y + bias


class _SoftplusD(torch.autograd.Function):
    """order-th derivative of softplus_beta (order 0 = the activation itself); its backward is the next derivative"""

    @staticmethod
    def forward(ctx, x, beta, order):
        x = _check(x)
        ctx.save_for_backward(x)
        ctx.beta, ctx.order = beta, order
        y = torch.empty_like(x)
        _lib.call('arcb200_softplus', x.data_ptr(), float(beta), int(order), y.data_ptr(), x.numel(), _lib.current_stream())
        return y

    @staticmethod
    def backward(ctx, g):
        (x,) = ctx.saved_tensors
        if ctx.order >= 2:
            raise RuntimeError('softplus: third-order derivatives are not provided')
        return g * _SoftplusD.apply(x, ctx.beta, ctx.order + 1), None, None


def softplus(x, beta=100.0):
    return _SoftplusD.apply(x, beta, 0)
