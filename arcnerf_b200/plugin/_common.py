"""Argument checks shared by the op wrappers: the reference's C++ layer enforces CUDA + contiguous +
dtype with TORCH_CHECK (arcnerf/ops/include/utils.h:17-27) and raises std::runtime_error on shape
mismatches (e.g. src/volume_func/volume_func.cpp:45-59); both surface as RuntimeError in Python."""
import torch


def check(t, name, dtype=None, shape=None):
    if not isinstance(t, torch.Tensor):
        raise RuntimeError('%s must be a tensor' % name)
    if not t.is_cuda:
        raise RuntimeError('%s must be a CUDA tensor' % name)
    if not t.is_contiguous():
        raise RuntimeError('%s must be a contiguous tensor' % name)
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError('%s must be a %s tensor' % (name, dtype))
    if shape is not None:
        if t.dim() != len(shape) or any(s is not None and s != d for s, d in zip(shape, t.shape)):
            raise RuntimeError('%s: expected shape %s, got %s' % (name, shape, tuple(t.shape)))
    return t


def f32c(t):
    """contiguous float32 view/copy (the wrappers of the reference call .contiguous() on rays)"""
    return t.contiguous() if t.dtype == torch.float32 else t.float().contiguous()
