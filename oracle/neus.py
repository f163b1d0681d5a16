"""CPU oracle of the NeuS-NGP foreground forward pass.  TEST INFRASTRUCTURE ONLY.

Composition, in the reference's order (arcnerf/models/neus_model.py:63-104, sdf_model.py:42-114), of pieces that are each
pinned to the reference: the C sampler oracle (oracle/sampler_oracle.c, bit-exact with the reference's host build), the
tiny-cuda-nn hash grid with dy/dx and the fused MLP (oracle/fields.py, pinned to tcnn run on B200), the SDF network as the
reference's own arithmetic (weight norm, softplus(beta), no bias: geo_rad_model/linear_network_module.py:174-197) in float64
numpy with its analytic input gradient, and sdf_to_alpha / ray_marching (oracle/render.py, pinned to the reference's functions).
The reference itself cannot run this configuration without a GPU (tinycudann), hence the composition."""
import numpy as np
import torch

from . import fields, render


def weight_norm(g, v):
    """nn.utils.weight_norm (dim 0): w = g * v / |v| per output row"""
    return g * v / np.linalg.norm(v, axis=1, keepdims=True)


def sdf_net(enc, denc_dx, w0, w1, beta=100.0):
    """enc (N,32) hash-grid features, denc_dx (N,32,3) -> sdf (N,), feat (N,16), normal (N,3) = d sdf / d x"""
    z = enc.astype(np.float64) @ w0.T.astype(np.float64)
    zb = z * beta
    h = np.where(zb > 20.0, z, np.log1p(np.exp(np.minimum(zb, 20.0))) / beta)
    s1 = np.where(zb > 20.0, 1.0, 1.0 / (1.0 + np.exp(-zb)))
    out = h @ w1.T.astype(np.float64)
    dsdf_denc = (s1 * w1[0][None, :].astype(np.float64)) @ w0.astype(np.float64)            # (N,32)
    normal = np.einsum('nk,nkd->nd', dsdf_denc, denc_dx.astype(np.float64))
    return out[:, 0], out[:, 1:], normal


def handle_mid_pts(zvals, mask_pts, n_sample):
    """neus_model.py:173-202 (masked branch), numpy"""
    sample_dist = (zvals[:, -1] - zvals[:, 0]) / np.float32(n_sample) * np.float32(0.5)
    final = zvals[:, -1] + sample_dist * np.float32(2.0)
    z2 = np.ones((zvals.shape[0], zvals.shape[1] + 1), np.float32) * final[:, None]
    m2 = np.concatenate([mask_pts, np.zeros_like(mask_pts[:, :1])], 1)
    z2[m2] = zvals[mask_pts]
    mid = np.float32(0.5) * (z2[:, 1:] + z2[:, :-1])
    mask_mid = np.concatenate([np.ones_like(mask_pts[:, :1]), mask_pts[:, :-1]], 1)
    return mid, z2, mask_mid


def forward(rays_o, rays_d, zvals, mask_pts, table, geom, aabb, w0, w1, rad_params, inv_s, speed_factor=10.0, n_sample=1024,
            white_bkg=True, bkg_color=None):
    """rays with at least one valid sample each.  Returns dict(rgb, depth, mask, normal) as numpy"""
    mid, z2, mask_mid = handle_mid_pts(zvals, mask_pts, n_sample)
    pts = rays_o[:, None, :] + rays_d[:, None, :] * mid[..., None]
    pv = pts[mask_mid].astype(np.float32)
    dv = np.broadcast_to(rays_d[:, None, :], pts.shape)[mask_mid].astype(np.float32)
    side = (aabb[1] - aabb[0]).astype(np.float32)
    unit = ((pv - aabb[0]) / side).astype(np.float32)
    enc, dydx = fields.grid_encode(unit, table.reshape(-1, 2), geom, 2, accumulate='fp32', with_dydx=True)
    dydx = dydx / side[None, None, :]                                            # d enc / d world position
    sdf_v, feat_v, nrm_v = sdf_net(enc, dydx, w0, w1)
    sh = fields.sh_encode(dv / np.linalg.norm(dv, axis=1, keepdims=True), 4, remap=True)
    x_in = np.concatenate([pv, sh, nrm_v.astype(np.float32), feat_v.astype(np.float32)], 1)          # 'pvnf': 3 + 16 + 3 + 16 = 38
    rad_v = fields.mlp_forward(x_in, rad_params, 38, 64, 2, 3, 'relu', 'sigmoid')
    n_rays, n_pts = mid.shape
    last = np.cumsum(mask_mid.sum(1)) - 1
    sdf = np.ones((n_rays, n_pts), np.float32) * sdf_v[last][:, None].astype(np.float32)
    rad = np.ones((n_rays, n_pts, 3), np.float32) * rad_v[last][:, None, :]
    nrm = np.ones((n_rays, n_pts, 3), np.float32) * nrm_v[last][:, None, :].astype(np.float32)
    sdf[mask_mid], rad[mask_mid], nrm[mask_mid] = sdf_v, rad_v, nrm_v
    slope = (rays_d[:, None, :] * nrm).sum(-1)
    iter_slope = -np.maximum(-slope, 0.0)                                         # inference: cos_anneal_ratio = 1
    s = float(np.exp(inv_s * speed_factor))
    alpha = render.sdf_to_alpha(torch.from_numpy(sdf), torch.from_numpy(z2), torch.from_numpy(iter_slope.astype(np.float32)), s)
    out = render.ray_marching(torch.from_numpy(sdf), torch.from_numpy(rad), torch.from_numpy(mid), False, None, white_bkg, alpha,
                              None if bkg_color is None else torch.from_numpy(bkg_color))
    nn_ = nrm / np.maximum(np.linalg.norm(nrm, axis=-1, keepdims=True), 1e-12)
    normal = (out['weights'].numpy()[..., None] * nn_).sum(-2)
    return dict(rgb=out['rgb'].numpy(), depth=out['depth'].numpy(), mask=out['mask'].numpy(), normal=normal, sdf=sdf, n_valid=int(mask_mid.sum()))
