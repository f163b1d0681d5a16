"""oracle/fields.py against the REFERENCE's own tiny-cuda-nn, executed on a B200 (tests/golden/tcnn_golden.npz, written by
tests/golden/make_golden_tcnn.py through oracle/tcnn_harness.cu).  This is what pins the hash-grid / SH / fused-MLP part of
the oracle: level geometry and hashing, table layout, per-corner fp16 accumulation, the loss-scale protocol of the backward
passes, the double backward, the flat (out x in) weight order of the MLP and its ones-padding.

Stated tolerances (and why they are not zero):
  grid forward       fp16-accumulate mode with the level scales the device computed: BIT-EXACT on every output; with the host's
                     libm scales (1 ulp off at 5 of 16 levels): <= 1 fp16 spacing at the table's magnitude (2^-12 for values
                     in [0.25, 0.5)); fp32-blend mode (what libarcb200 computes): <= 2 such spacings
  grid gradients     tcnn accumulates dL/dparams with half2 atomics at loss scale 128: rel 2e-3 of the largest entry
  SH                 fp16 output rounding: 1 fp16 ulp of 1.0 (4.9e-4)
  MLP forward        tcnn accumulates in fp16 inside wmma, the oracle exactly: abs 3e-3
  MLP backward       on rows whose ReLU gates are 0.02 away from switching: dL/dx and dL/dparams rel 1e-2 of the largest entry
"""
import os

import numpy as np
import pytest

import tcnn_cases as tc
from oracle import fields

GOLD = os.path.join(os.path.dirname(__file__), 'golden', 'tcnn_golden.npz')


@pytest.fixture(scope='module')
def G():
    return np.load(GOLD)


@pytest.fixture(scope='module')
def geom():
    g = tc.GRID
    return fields.grid_geometry(g['n_levels'], g['log2_hashmap_size'], g['base_res'], tc.per_level_scale())


def rel(a, b):
    return float(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64)).max() / np.abs(b).max())


def ulps16(a, b, magnitude=0.3):
    """largest difference in units of the fp16 spacing at `magnitude` (default: the size of the test table's entries, whose
    rounded partial sums set the error of an interpolated feature)"""
    spacing = float(np.spacing(np.float16(magnitude)))
    return np.abs(np.asarray(a, np.float32) - np.asarray(b, np.float32)) / spacing


def test_grid_geometry_matches_tcnn(G, geom):
    assert int(geom[0][-1]) * tc.GRID['n_features'] == int(G['grid_n_params']) == 12196240
    assert abs(float(G['grid_per_level_scale']) - geom[4]) < 1e-12
    # the kernels recompute the scale per thread with the device's exp2f: at most an ulp away from the host value
    dev = G['grid_device_scales']
    assert np.all(np.abs(dev - geom[1]) <= np.spacing(geom[1]) * 1.01)
    # and the resolutions that follow from either are the same (the hard case is level 15: 2047.0 -> 2048)
    assert np.array_equal(np.ceil(dev).astype(np.uint32) + 1, geom[2])


def test_grid_forward_matches_tcnn(G, geom):
    table, x = tc.grid_table(int(G['grid_n_params'])).reshape(-1, 2), tc.grid_points()
    ref = G['grid_fwd'].astype(np.float32)
    assert np.array_equal(G['grid_fwd'], G['grid_inference'])
    # tcnn's own arithmetic (fp16 rounding after every corner), with the scales its kernels used
    dev_geom = (geom[0], G['grid_device_scales'], geom[2], geom[3], geom[4])
    y = fields.grid_encode(x, table, dev_geom, 2, accumulate='fp16')
    assert np.array_equal(y, ref)                      # bit-exact: the oracle IS tcnn's arithmetic
    # host scales: within one fp16 spacing
    y = fields.grid_encode(x, table, geom, 2, accumulate='fp16')
    assert ulps16(y, ref).max() <= 1.0
    # the fp32 blend libarcb200 computes
    y32 = fields.grid_encode(x, table, geom, 2, accumulate='fp32')
    assert ulps16(y32, ref).max() <= 2.0


def test_grid_backward_matches_tcnn(G, geom):
    n_params = int(G['grid_n_params'])
    table, x = tc.grid_table(n_params).reshape(-1, 2), tc.grid_points()
    dy = tc.grid_dy().astype(np.float32)
    _, dydx = fields.grid_encode(x, table, geom, 2, accumulate='fp32', with_dydx=True)
    dLdx = np.einsum('nk,nkd->nd', dy.astype(np.float64), dydx.astype(np.float64))
    assert rel(dLdx, G['grid_dLdx']) < 1e-4
    gt = fields.grid_backward_table(x, dy, geom, 2).reshape(-1)
    ref = np.zeros(n_params, np.float64)
    ref[G['grid_dLdp_idx']] = G['grid_dLdp_val']
    assert rel(gt, ref) < 2e-3
    # the same table entries are touched (entries whose fp16 gradient rounds to zero aside)
    touched, ref_touched = np.flatnonzero(gt), G['grid_dLdp_idx']
    assert np.isin(ref_touched, touched).all() and touched.size - ref_touched.size < 1e-3 * touched.size


def test_grid_double_backward_matches_tcnn(G, geom):
    n_params = int(G['grid_n_params'])
    table, x = tc.grid_table(n_params).reshape(-1, 2), tc.grid_points()
    dy, v = tc.grid_dy().astype(np.float32), tc.grid_v()
    g_tab, d_dy, d_x = fields.grid_double_backward(x, table, dy, v, geom, 2)
    ref = np.zeros(n_params, np.float64)
    ref[G['grid_bb_dLdp_idx']] = G['grid_bb_dLdp_val']
    assert np.isfinite(ref).all()
    assert rel(g_tab.reshape(-1), ref) < 2e-3
    assert rel(d_dy, G['grid_bb_dLddLdy']) < 2e-3          # fp16 output
    assert rel(d_x, G['grid_bb_dLdx']) < 1e-4


def test_sh_matches_tcnn(G):
    sh = fields.sh_encode(tc.sh_dirs(), 4, remap=True)
    assert np.abs(sh - G['sh_fwd'].astype(np.float32)).max() <= 4.9e-4


@pytest.mark.parametrize('k', range(len(tc.MLPS)))
def test_mlp_matches_tcnn(G, k):
    name, i, w, h, o, act, oact = tc.MLPS[k]
    p = tc.mlp_params(i, w, h, o, seed=200 + k)
    assert int(G['mlp_%s_n_out' % name]) == 16                     # outputs padded to 16 columns
    assert np.array_equal(G['mlp_%s_fwd' % name], G['mlp_%s_inference' % name])
    y = fields.mlp_forward(tc.mlp_inputs(i, seed=300 + k), p, i, w, h, o, act, oact)
    assert np.abs(y - G['mlp_%s_fwd' % name].astype(np.float32)[:, :o]).max() < 3e-3
    xs = tc.mlp_inputs_gate_stable(i, w, h, o, p, seed=500 + k)
    ys = fields.mlp_forward(xs, p, i, w, h, o, act, oact)
    assert np.abs(ys - G['mlp_%s_stable_fwd' % name].astype(np.float32)[:, :o]).max() < 3e-3
    gy = tc.mlp_dy(16, o, seed=400 + k).astype(np.float32)
    dx, dw = fields.mlp_backward(xs, p, gy[:, :o], i, w, h, o, act, oact)
    assert rel(dx, G['mlp_%s_dLdx' % name]) < 1e-2
    assert rel(dw, G['mlp_%s_dLdp' % name]) < 1e-2


def test_initial_parameter_distributions(G):
    """tcnn draws its initial parameters from pcg32(1337); this library and the oracle draw from torch's generator.  What is
    pinned is the distribution: hash table U(-1e-4, 1e-4) (grid.h:1353-1356), MLP matrices xavier-uniform."""
    lo, hi, mean, std = G['grid_init_stats']
    assert -1.0001e-4 <= lo < -0.99e-4 and 0.99e-4 < hi <= 1.0001e-4 and abs(mean) < 1e-7
    assert abs(std - 2e-4 / np.sqrt(12.0)) < 1e-7
    for k, (name, i, w, h, o, _, _) in enumerate(tc.MLPS):
        init = G['mlp_%s_init' % name]
        shapes, _, _ = fields.mlp_shapes(i, w, h, o)
        off = 0
        for (no, ni) in shapes:
            m = init[off:off + no * ni]
            bound = np.sqrt(6.0 / (ni + no))
            assert np.abs(m).max() <= bound * 1.0001 and np.abs(m).max() > 0.97 * bound, (name, no, ni)
            off += no * ni
