"""Seeded rays and weight seeds of the model-level goldens (tests/golden/models_golden.npz)."""
import numpy as np

N_RAYS = 96
SEEDS = dict(fg_coarse=10, fg_fine=11, bkg_coarse=12, bkg_fine=13)

# the `model` sections of configs/models/nerf.yaml and nerfpp.yaml (+ background), with jitter and density noise off
GEO = dict(W=256, D=8, skips=[4], encoder=dict(type='FreqEmbedder', input_dim=3, n_freqs=10), W_feat=256, geometric_init=False)
RAD = dict(mode='vf', W=128, D=1, encoder=dict(view=dict(type='FreqEmbedder', input_dim=3, n_freqs=4)), W_feat_in=256)
NERF_CFG = dict(rays=dict(bounding_radius=3.0, near=1.5, far=8.0, n_sample=64, n_importance=128, inverse_linear=False, perturb=False,
                          add_inf_z=True, noise_std=0.0, white_bkg=False), chunk_rays=32768, geometry=GEO, radiance=RAD)
NERFPP_FG_CFG = dict(rays=dict(bounding_radius=3.0, near=1.5, n_sample=32, n_importance=64, inverse_linear=False, perturb=False,
                               add_inf_z=False, noise_std=0.0, white_bkg=False), chunk_rays=32768, geometry=GEO, radiance=RAD)
NERFPP_BKG_CFG = dict(rays=dict(bounding_radius=3.0, n_sample=32, n_importance=64, perturb=False, add_inf_z=True, noise_std=0.0,
                                white_bkg=False), chunk_rays=32768,
                      geometry=dict(GEO, encoder=dict(type='FreqEmbedder', input_dim=4, n_freqs=10)), radiance=RAD)


def rays(n=N_RAYS, seed=3):
    """cameras on a sphere of radius 2 (inside the bounding radius 3) looking roughly at the origin"""
    rs = np.random.RandomState(seed)
    o = rs.normal(size=(n, 3))
    o = o / np.linalg.norm(o, axis=1, keepdims=True) * 2.0
    d = -o / 2.0 + 0.25 * rs.normal(size=(n, 3))
    d = d / np.linalg.norm(d, axis=1, keepdims=True)
    return o.astype(np.float32), d.astype(np.float32)
