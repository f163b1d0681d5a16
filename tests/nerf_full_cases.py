"""Seeded weights and inputs of the full-width vanilla-NeRF network goldens (tests/golden/nerf_full_golden.npz): shared by
the generator (which loads them into the REFERENCE's own GeoNet / RadianceNet) and by the GPU test."""
import numpy as np

N_PTS = 700            # not a multiple of the 256-row tile pair: exercises the ragged tail
CASES = {
    # name: (d_x, n_freq_x, n_freq_v, W, D, skips, W_rad)
    'nerf': (3, 10, 4, 256, 8, [4], 128),          # configs/models/nerf.yaml
    'nerfpp_bkg': (4, 10, 4, 256, 8, [4], 128),    # configs/models/nerfpp.yaml background (x/r, y/r, z/r, 1/r)
}


def layer_shapes(d_x, n_freq_x, n_freq_v, W, D, skips, W_rad):
    ex, ev = d_x * (2 * n_freq_x + 1), 3 * (2 * n_freq_v + 1)
    geo = []
    for i in range(D + 1):
        fan_in = ex if i == 0 else (W + ex if (i - 1) in skips else W)
        geo.append((1 + W if i == D else W, fan_in))
    return geo, [(W_rad, ev + W), (3, W_rad)]


def weights(name, seed=0):
    """-> (geo_state, rad_state) numpy dicts in the reference's state-dict naming; Linear-default bounds times 1.3"""
    rs = np.random.RandomState(seed)
    geo_shapes, rad_shapes = layer_shapes(*CASES[name])
    out = []
    for shapes in (geo_shapes, rad_shapes):
        sd = {}
        for i, (o, fan_in) in enumerate(shapes):
            b = 1.3 / np.sqrt(fan_in)
            sd['layers.%d.weight' % i] = rs.uniform(-b, b, (o, fan_in)).astype(np.float32)
            sd['layers.%d.bias' % i] = rs.uniform(-b, b, (o,)).astype(np.float32)
        out.append(sd)
    return out


def inputs(name, seed=1):
    d_x = CASES[name][0]
    rs = np.random.RandomState(seed)
    x = rs.uniform(-1.0, 1.0, (N_PTS, d_x)).astype(np.float32)
    dirs = (rs.normal(size=(N_PTS, 3)) * rs.uniform(0.5, 2.0, (N_PTS, 1))).astype(np.float32)     # not normalised on purpose
    w_sigma = rs.normal(size=(N_PTS, 1)).astype(np.float32)
    w_rgb = rs.normal(size=(N_PTS, 3)).astype(np.float32)
    return x, dirs, w_sigma, w_rgb
