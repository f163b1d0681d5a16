"""The CUDA path against the REFERENCE's own tiny-cuda-nn outputs (tests/golden/tcnn_golden.npz, produced on a B200 by
tests/golden/make_golden_tcnn.py): hash-grid forward / input gradient / table gradient / double backward, SH, and the fused
MLPs of the NGP and NeuS-NGP configurations forward and backward, all through the C ABI.  Tolerances as stated in
tests/test_oracle_tcnn.py (tcnn rounds to fp16 after every corner and accumulates the MLP in fp16; libarcb200 blends and
accumulates in fp32)."""
import os

import numpy as np
import pytest
import torch

import tcnn_cases as tc
from test_oracle_tcnn import rel, ulps16

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), 'golden', 'tcnn_golden.npz')
cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.fixture(scope='module')
def G():
    return np.load(GOLD)


@pytest.fixture(scope='module')
def grid():
    from arcnerf_b200 import functional as F
    g = tc.GRID
    geom = F.GridGeometry(g['n_levels'], g['n_features'], g['log2_hashmap_size'], g['base_res'], tc.per_level_scale())
    table = cu(tc.grid_table(geom.n_params))
    return geom, F.cast_f16(table)


def test_grid_forward_and_input_gradient(G, grid):
    from arcnerf_b200 import functional as F
    geom, table_h = grid
    assert geom.n_params == int(G['grid_n_params'])
    x = cu(tc.grid_points())
    y, dydx = F.grid_forward(x, table_h, geom, want_dydx=True)
    assert ulps16(y.float().cpu().numpy(), G['grid_fwd'].astype(np.float32)).max() <= 2.0
    dy = cu(tc.grid_dy())
    dLdx = F.grid_backward_input(dy, dydx, geom)
    assert rel(dLdx.cpu().numpy(), G['grid_dLdx']) < 1e-4


def test_grid_table_gradient(G, grid):
    from arcnerf_b200 import functional as F
    geom, _ = grid
    gt = F.grid_backward(cu(tc.grid_points()), cu(tc.grid_dy()), geom).cpu().numpy()
    ref = np.zeros(geom.n_params, np.float64)
    ref[G['grid_dLdp_idx']] = G['grid_dLdp_val']
    assert rel(gt, ref) < 2e-3


def test_grid_double_backward(G, grid):
    from arcnerf_b200 import functional as F
    geom, table_h = grid
    gt = torch.zeros(geom.n_params, dtype=torch.float32, device='cuda')
    _, d_dy, d_x = F.grid_double_backward(cu(tc.grid_points()), table_h, cu(tc.grid_dy()), cu(tc.grid_v()), geom, grad_table=gt)
    ref = np.zeros(geom.n_params, np.float64)
    ref[G['grid_bb_dLdp_idx']] = G['grid_bb_dLdp_val']
    assert rel(gt.cpu().numpy(), ref) < 2e-3
    assert rel(d_dy.cpu().numpy(), G['grid_bb_dLddLdy']) < 2e-3
    assert rel(d_x.cpu().numpy(), G['grid_bb_dLdx']) < 1e-4


def test_sh(G):
    from arcnerf_b200 import functional as F
    sh = F.sh_forward(cu(tc.sh_dirs()), 4, remap=True)
    assert np.abs(sh.float().cpu().numpy() - G['sh_fwd'].astype(np.float32)).max() <= 4.9e-4


@pytest.mark.parametrize('k', range(len(tc.MLPS)))
def test_mlp_forward_and_backward(G, k):
    from arcnerf_b200 import functional as F
    name, i, w, h, o, act, oact = tc.MLPS[k]
    p = tc.mlp_params(i, w, h, o, seed=200 + k)
    spec = F.MlpSpec(i, w, h, o, act, oact)
    assert spec.n_params == p.size
    blob = F.mlp_pack(cu(p), spec)
    y = F.mlp_forward(cu(tc.mlp_inputs(i, seed=300 + k)), blob, spec)
    assert np.abs(y.cpu().numpy() - G['mlp_%s_fwd' % name].astype(np.float32)[:, :o]).max() < 3e-3
    xs = cu(tc.mlp_inputs_gate_stable(i, w, h, o, p, seed=500 + k))
    gy = cu(tc.mlp_dy(16, o, seed=400 + k).astype(np.float32)[:, :o])
    dx, dw = F.mlp_backward(xs, blob, spec, gy)
    assert rel(dx.cpu().numpy(), G['mlp_%s_dLdx' % name]) < 1e-2
    assert rel(dw.cpu().numpy(), G['mlp_%s_dLdp' % name]) < 1e-2
