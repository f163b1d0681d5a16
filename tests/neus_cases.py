"""The vanilla (frequency-embedded, sphere-bounded) NeuS configuration used for the up-sampling golden: configs/models/neus.yaml
with narrower networks and jitter off."""
import numpy as np

CFG = dict(
    obj_bound=dict(sphere=dict(radius=1.5)),
    rays=dict(radius_bound=1.5, n_sample=32, n_importance=32, n_iter=4, inverse_linear=False, perturb=False, add_inf_z=False,
              noise_std=0.0, white_bkg=False),
    params=dict(speed_factor=10, init_var=0.05, anneal_end=50000), chunk_rays=4096, chunk_pts=131072,
    geometry=dict(W=64, D=4, skips=[2], encoder=dict(type='FreqEmbedder', input_dim=3, n_freqs=6), W_feat=32, geometric_init=True,
                  radius_init=0.75, weight_norm=True, skip_reduce_output=True, norm_skip=True, act_cfg=dict(type='softplus', beta=100)),
    radiance=dict(mode='pvnf', W=64, D=2, encoder=dict(pts=dict(type='FreqEmbedder', input_dim=3, n_freqs=0),
                                                       view=dict(type='FreqEmbedder', input_dim=3, n_freqs=4)),
                  W_feat_in=32, weight_norm=True))
N_RAYS = 96


def rays(seed=2, n=N_RAYS):
    rs = np.random.RandomState(seed)
    o = rs.normal(size=(n, 3))
    o = o / np.linalg.norm(o, axis=1, keepdims=True) * 3.0
    d = -o / 3.0 + 0.2 * rs.normal(size=(n, 3))
    d = d / np.linalg.norm(d, axis=1, keepdims=True)
    return o.astype(np.float32), d.astype(np.float32)
