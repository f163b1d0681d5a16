# -*- coding: utf-8 -*-
"""Drop-in for arcnerf/models/base_modules/encoding/sh_encoder.py (backend b200 / tcnn-compatible): spherical
harmonics embedding of view directions, degree 1..5, evaluated by libarcb200 (sh_encoder.py:20-123)."""
import torch
import torch.nn as nn

from .. import functional as F
from .hashgrid_encoder import cast_torch_dtype
from .registry import ENCODER_REGISTRY


@ENCODER_REGISTRY.register()
class SHEmbedder(nn.Module):
    def __init__(self, input_dim=3, n_freqs=4, include_input=True, backend=None, dtype='torch.float16', *args, **kwargs):
        super(SHEmbedder, self).__init__()
        assert input_dim == 3, 'SHEmbedder should has input_dim==3...'
        assert 1 <= n_freqs <= 5, 'Should have degree 1~5 for encoding...'
        if backend is None:
            backend = 'b200'
        assert backend in ['b200', 'tcnn'], 'Invalid backend: this build provides the CUDA (b200/tcnn-compatible) path only'
        self.input_dim, self.n_freqs, self.include_input, self.backend = input_dim, n_freqs, include_input, backend
        self.dtype = cast_torch_dtype(dtype)
        self.out_dim = n_freqs ** 2 + include_input * input_dim

    def get_output_dim(self):
        return self.out_dim

    def forward(self, xyz):
        """xyz (B,3) direction; no gradient flows to the direction (the reference's view dirs carry none either)"""
        assert len(xyz.shape) == 2 and xyz.shape[-1] == 3, 'Must be (B, 3) direction'
        with torch.no_grad():
            sh = F.sh_forward(xyz.detach().contiguous().float(), self.n_freqs, remap=True,
                              out_dtype=torch.float16).to(self.dtype)
        # include_input is False in every NGP / NeuS-NGP configuration (configs/models/nerf_ngp.yaml:56); the concatenation of the
        # raw direction is the one torch op of this module and stays off the hot path
        return torch.cat([xyz, sh], dim=-1) if self.include_input else sh
