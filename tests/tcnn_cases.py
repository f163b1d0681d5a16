"""Seeded inputs of the tiny-cuda-nn golden vectors (tests/golden/tcnn_golden.npz).  Shared by the generator
(tests/golden/make_golden_tcnn.py, which runs the reference's own tiny-cuda-nn on a B200) and by the tests that
check oracle/fields.py and the CUDA path against the file.  Everything is regenerated from numpy's legacy
RandomState streams (frozen by numpy's compatibility policy), so only the reference's OUTPUTS are stored.
"""
import math

import numpy as np

N_GRID_PTS = 2048          # multiple of tcnn's batch granularity (128)
N_MLP_ROWS = 1024

# HashGridEmbedder as ArcNerf configures tcnn for NGP / NeuS-NGP (hashgrid_encoder.py:102-113)
GRID = dict(n_levels=16, n_features=2, log2_hashmap_size=19, base_res=16, max_res=2048)

# (name, in_dim, width, n_hidden, out_dim, hidden activation, output activation) -- tcnn_fusedmlp_module.py:66-77,162-173
MLPS = [
    ('geo', 32, 64, 1, 16, 'ReLU', 'None'),            # FusedMLPGeoNet of nerf_ngp.yaml
    ('rad', 32, 64, 2, 3, 'ReLU', 'Sigmoid'),          # FusedMLPRadianceNet, mode 'fv'
    ('rad_pvnf', 38, 64, 2, 3, 'ReLU', 'Sigmoid'),     # NeuS-NGP radiance head: 38 inputs padded to 48 with ones
]


def per_level_scale():
    """exp(log(max_res/base_res)/(L-1)) evaluated in float32 like hashgrid_encoder.py:81"""
    import torch
    return float(torch.exp(torch.log(torch.tensor(GRID['max_res'] / GRID['base_res'])) / (float(GRID['n_levels']) - 1)))


def grid_table(n_params, seed=101, scale=0.3):
    return (np.random.RandomState(seed).uniform(-1.0, 1.0, n_params) * scale).astype(np.float32)


def grid_points(n=N_GRID_PTS, seed=102):
    """uniform points in [0,1)^3 plus edge cases: box corners, exact cell boundaries of several levels"""
    x = np.random.RandomState(seed).uniform(0.0, 1.0, (n, 3)).astype(np.float32)
    x[0] = (0.0, 0.0, 0.0)
    x[1] = (1.0, 1.0, 1.0)
    x[2] = (0.5, 0.5, 0.5)
    x[3] = (0.0, 1.0, 0.5)
    x[4] = (0.17, 0.55, 0.79)                # tcnn/scripts/test_grid_bwdbwd.py:106-140
    for i, r in enumerate((16, 23, 59, 295, 2048)):
        x[5 + i] = (3.0 / r, 7.0 / r, (r - 1.0) / r)
    return x


def grid_dy(n=N_GRID_PTS, dim=32, seed=103):
    return np.random.RandomState(seed).normal(0.0, 1.0, (n, dim)).astype(np.float16)


def grid_v(n=N_GRID_PTS, seed=104):
    """dL/d(dL/dx) of the double backward; small, so that tcnn's fp16 table gradient (x loss scale 128 x level scale up to
    2047) stays below the fp16 maximum"""
    return (np.random.RandomState(seed).normal(0.0, 1.0, (n, 3)) * 1e-3).astype(np.float32)


def sh_dirs(n=N_GRID_PTS, seed=105):
    d = np.random.RandomState(seed).normal(0.0, 1.0, (n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[0] = (0.0, 0.0, 1.0)
    d[1] = (1.0, 0.0, 0.0)
    d[2] = (0.0, -1.0, 0.0)
    return d.astype(np.float32)


def mlp_n_params(in_dim, width, n_hidden, out_dim):
    in_pad, out_pad = (in_dim + 15) // 16 * 16, (out_dim + 15) // 16 * 16
    return width * in_pad + (n_hidden - 1) * width * width + out_pad * width


def mlp_params(in_dim, width, n_hidden, out_dim, seed):
    """xavier-uniform per matrix, tcnn layout (input layer first, (out x in) row-major, last matrix padded to 16 rows)"""
    rs = np.random.RandomState(seed)
    in_pad, out_pad = (in_dim + 15) // 16 * 16, (out_dim + 15) // 16 * 16
    shapes = [(width, in_pad)] + [(width, width)] * (n_hidden - 1) + [(out_pad, width)]
    return np.concatenate([rs.uniform(-1.0, 1.0, o * i) * math.sqrt(6.0 / (i + o)) for o, i in shapes]).astype(np.float32)


def mlp_inputs(in_dim, n=N_MLP_ROWS, seed=0):
    return np.random.RandomState(seed).uniform(-1.0, 1.0, (n, in_dim)).astype(np.float32)


def mlp_inputs_gate_stable(in_dim, width, n_hidden, out_dim, params, n=N_MLP_ROWS, seed=0, margin=0.02):
    """Rows whose hidden pre-activations all stay `margin` away from zero: tcnn accumulates in fp16, so a ReLU whose argument
    is within its rounding error of zero can be open in one implementation and shut in the other, which changes whole rows of
    the gradients.  Backward parity is checked on rows where that cannot happen (the forward is checked on unfiltered rows).
    Plain float64 forward on fp16-rounded weights and inputs padded with ones, no library involved."""
    rs = np.random.RandomState(seed)
    in_pad = (in_dim + 15) // 16 * 16
    shapes = [(width, in_pad)] + [(width, width)] * (n_hidden - 1)
    mats, off = [], 0
    for o, i in shapes:
        mats.append(params[off:off + o * i].reshape(o, i).astype(np.float16).astype(np.float64))
        off += o * i
    rows = []
    while sum(len(r) for r in rows) < n:
        x = rs.uniform(-1.0, 1.0, (8 * n, in_dim)).astype(np.float32)
        h = np.ones((x.shape[0], in_pad))
        h[:, :in_dim] = x.astype(np.float16).astype(np.float64)
        ok = np.ones(x.shape[0], bool)
        for w in mats:
            z = h @ w.T
            ok &= np.abs(z).min(1) > margin
            h = np.maximum(z, 0).astype(np.float16).astype(np.float64)
        rows.append(x[ok])
    return np.concatenate(rows)[:n]


def mlp_dy(out_pad, out_dim, n=N_MLP_ROWS, seed=0):
    """gradient w.r.t. the padded fp16 output; pad columns carry no gradient (the reference slices them off)"""
    g = np.zeros((n, out_pad), np.float16)
    g[:, :out_dim] = np.random.RandomState(seed).normal(0.0, 1.0, (n, out_dim)).astype(np.float16)
    return g
