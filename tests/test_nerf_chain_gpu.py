"""The vanilla-NeRF network pair as one tensor-core chain (csrc/dense_chain.cu, plugin/nerf_mlp.py) against the REFERENCE's
own GeoNet + RadianceNet at full width (tests/golden/nerf_full_golden.npz, made by tests/golden/make_golden_nerf_full.py from
the reference's modules with the seeded weights of tests/nerf_full_cases.py).
Tolerances: the reference computes in fp32; the chain uses fp16 operands (weights, embeddings, activations) with fp32
accumulation over 9 + 2 layers -> sigma abs 3e-2 (values of order 1), rgb abs 4e-3, gradients 3e-2 of the largest entry."""
import os

import numpy as np
import pytest
import torch

import nerf_full_cases as nc

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'nerf_full_golden.npz'))


def build(name):
    from arcnerf_b200.plugin.nerf_mlp import FusedNeRFNets
    d_x, nfx, nfv, W, D, skips, W_rad = nc.CASES[name]
    nets = FusedNeRFNets(d_x, nfx, nfv, W, D, skips, W_rad)
    gs, rs = nc.weights(name)
    nets.load_reference_state(gs, rs)
    return nets


@pytest.mark.parametrize('name', list(nc.CASES))
def test_forward_matches_reference_modules(name):
    nets = build(name)
    x, dirs, _, _ = (torch.from_numpy(a).cuda() for a in nc.inputs(name))
    with torch.no_grad():
        sigma, rgb = nets(x, dirs)
    torch.cuda.synchronize()
    assert sigma.shape == (nc.N_PTS, 1) and rgb.shape == (nc.N_PTS, 3)
    ref_s, ref_c = G[name + '/sigma'], G[name + '/rgb']
    assert np.abs(sigma.cpu().numpy() - ref_s).max() < 3e-2 * max(1.0, np.abs(ref_s).max())
    assert np.abs(rgb.cpu().numpy() - ref_c).max() < 4e-3


@pytest.mark.parametrize('name', list(nc.CASES))
def test_parameter_gradients_match_reference_autograd(name):
    nets = build(name)
    x, dirs, w_s, w_c = (torch.from_numpy(a).cuda() for a in nc.inputs(name))
    sigma, rgb = nets(x, dirs)
    ((sigma * w_s).sum() + (rgb * w_c).sum()).backward()
    torch.cuda.synchronize()
    g = nets.flat.grad
    off = 0
    n_geo = nets.arch[4] + 1
    for li, (o, i) in enumerate(nets.shapes):
        tag, k = ('geo', li) if li < n_geo else ('rad', li - n_geo)
        gw = g[off:off + o * i].view(o, i).cpu().numpy()
        gb = g[off + o * i:off + o * i + o].cpu().numpy()
        off += o * i + o
        key = '%s/grad/%s/layers.%d' % (name, tag, k)
        scale = float(G[key + '.weight.absmax'])
        assert np.abs(gw[:8] - G[key + '.weight.rows8']).max() < 3e-2 * scale, key
        ref_cs = G[key + '.weight.colsum']
        assert np.abs(gw.sum(0) - ref_cs).max() < 3e-2 * max(np.abs(ref_cs).max(), scale), key
        ref_b = G[key + '.bias']
        assert np.abs(gb - ref_b).max() < 3e-2 * np.abs(ref_b).max(), key


def test_state_dict_round_trip_and_repack_after_update():
    nets = build('nerf')
    geo, rad = nets.reference_state()
    gs, rs = nc.weights('nerf')
    assert all(np.array_equal(geo[k].cpu().numpy(), gs[k]) for k in gs) and all(np.array_equal(rad[k].cpu().numpy(), rs[k]) for k in rs)
    x, dirs, _, _ = (torch.from_numpy(a).cuda() for a in nc.inputs('nerf'))
    with torch.no_grad():
        s0, _ = nets(x, dirs)
        nets.flat.mul_(0.5)                       # an optimiser step changes the parameters in place: blobs are re-packed
        s1, _ = nets(x, dirs)
    assert float((s0 - s1).abs().max()) > 1e-3


def test_unsupported_architecture_is_an_error():
    from arcnerf_b200.plugin.nerf_mlp import FusedNeRFNets
    with pytest.raises(RuntimeError):
        FusedNeRFNets(3, 10, 4, 128, 8, [4], 128)
