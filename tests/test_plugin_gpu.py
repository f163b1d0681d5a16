"""GPU: the nn.Module / function mirrors of the reference's plugin surface (arcnerf_b200/plugin) - same names,
arguments, returned shapes and state_dict keys, autograd (incl. the grid's double backward) through the C ABI."""
import os

import numpy as np
import pytest
import torch

import render_cases
from arcnerf_b200.plugin import ray_helper
from arcnerf_b200.plugin.hashgrid_encoder import HashGridEmbedder
from arcnerf_b200.plugin.registry import ENCODER_REGISTRY, MODULE_REGISTRY
from arcnerf_b200.plugin.sh_encoder import SHEmbedder
from arcnerf_b200.plugin.tcnn_fusedmlp_module import FusedMLPGeoNet, FusedMLPRadianceNet
from arcnerf_b200.plugin.trunc_exp import TruncExp
from oracle import fields, render

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'render_golden.npz'))

# configs/models/nerf_ngp.yaml (geometry / radiance blocks)
GEO_CFG = dict(W=64, D=1, W_feat=16, out_act_cfg=dict(type='TruncExp'),
               encoder=dict(type='HashGridEmbedder', input_dim=3, include_input=False, side=2.0, n_levels=16,
                            n_feat_per_entry=2, hashmap_size=19, base_res=16, max_res=2048, backend='tcnn'))
RAD_CFG = dict(mode='fv', W=64, D=2, W_feat_in=16,
               encoder=dict(view=dict(type='SHEmbedder', include_input=False, input_dim=3, n_freqs=4, backend='tcnn')))


def test_registry_and_state_dict_keys(cuda):
    assert 'HashGridEmbedder' in ENCODER_REGISTRY and 'SHEmbedder' in ENCODER_REGISTRY
    assert 'FusedMLPGeoNet' in MODULE_REGISTRY and 'FusedMLPRadianceNet' in MODULE_REGISTRY
    geo = MODULE_REGISTRY.get('FusedMLPGeoNet')(**GEO_CFG).to(cuda)
    rad = MODULE_REGISTRY.get('FusedMLPRadianceNet')(**RAD_CFG).to(cuda)
    sd = geo.state_dict()
    assert sd['embed_fn.hashgrid_encode_tcnn.params'].shape == (12196240,)      # tcnn flat table, SURVEY.md §2.2
    assert sd['layers.params'].shape == (3072,) and rad.state_dict()['layers.params'].shape == (7168,)
    assert geo.embed_fn.get_output_dim() == 32 and rad.init_input_dim == 32


def test_ngp_heads_forward_backward_vs_oracle(cuda):
    torch.manual_seed(0)
    geo = FusedMLPGeoNet(**GEO_CFG).to(cuda)
    rad = FusedMLPRadianceNet(**RAD_CFG).to(cuda)
    with torch.no_grad():
        geo.embed_fn.hashgrid_encode_tcnn.params.uniform_(-0.3, 0.3)
    n = 3000
    rng = np.random.RandomState(0)
    x = rng.uniform(-1, 1, (n, 3)).astype(np.float32)
    d = render_cases.unit_dirs(n, 1) * 1.7                      # not normalised: the net normalises
    xt = torch.from_numpy(x).to(cuda)
    sigma, feat = geo(xt)
    rgb = rad(xt, torch.from_numpy(d).to(cuda), None, feat)
    assert sigma.shape == (n, 1) and feat.shape == (n, 16) and rgb.shape == (n, 3)
    assert sigma.dtype == torch.float32 and rgb.dtype == torch.float32
    assert float(rgb.min()) >= 0.0 and float(rgb.max()) <= 1.0
    # oracle
    emb = geo.embed_fn
    og = fields.grid_geometry(16, 19, 16, float(emb.per_level_scale))
    table = emb.hashgrid_encode_tcnn.params.detach().cpu().numpy()
    unit = ((xt - emb.min_xyz) / (emb.max_xyz - emb.min_xyz)).cpu().numpy()
    enc = fields.grid_encode(unit, table.reshape(-1, 2), og, 2)
    gp, rp = geo.layers.params.detach().cpu().numpy(), rad.layers.params.detach().cpu().numpy()
    go = fields.mlp_forward(enc, gp, 32, 64, 1, 16).astype(np.float16).astype(np.float32)
    np.testing.assert_allclose(feat.cpu().detach().numpy(), go, atol=4e-3, rtol=4e-3)
    np.testing.assert_allclose(sigma[:, 0].cpu().detach().numpy(), np.exp(go[:, 0]), rtol=1e-2)
    dn = d / np.linalg.norm(d, axis=1, keepdims=True)
    sh = fields.sh_encode(dn, 4, remap=True).astype(np.float16).astype(np.float32)
    ro = fields.mlp_forward(np.concatenate([go, sh], 1), rp, 32, 64, 2, 3, 'relu', 'sigmoid')
    np.testing.assert_allclose(rgb.cpu().detach().numpy(), ro, atol=4e-3)
    # backward reaches all three parameter tensors
    (rgb.sum() + sigma.sum() * 1e-3).backward()
    for p in (geo.embed_fn.hashgrid_encode_tcnn.params, geo.layers.params, rad.layers.params):
        assert p.grad is not None and torch.isfinite(p.grad).all() and float(p.grad.abs().max()) > 0
    # parameter update invalidates the cached fp16 copies
    with torch.no_grad():
        geo.layers.params.mul_(0.5)
        geo.embed_fn.hashgrid_encode_tcnn.params.mul_(2.0)
    s2, _ = geo(xt)
    assert not torch.allclose(s2, sigma)


def test_hashgrid_input_gradient_and_eikonal_double_backward(cuda):
    """NeuS path: normals = d sdf / dx by autograd with create_graph, then an eikonal loss on them (second order)."""
    emb = HashGridEmbedder(n_levels=6, n_feat_per_entry=2, hashmap_size=10, base_res=4, max_res=64, side=2.0,
                           include_input=True, dtype='torch.float32', backend='tcnn').to(cuda)
    assert emb.get_output_dim() == 15
    with torch.no_grad():
        emb.hashgrid_encode_tcnn.params.uniform_(-1, 1)
    lin = torch.nn.Linear(15, 1).to(cuda)
    rng = np.random.RandomState(4)
    x = torch.from_numpy(rng.uniform(-0.9, 0.9, (500, 3)).astype(np.float32)).to(cuda).requires_grad_(True)
    sdf = lin(emb(x))
    normal = torch.autograd.grad(sdf, x, torch.ones_like(sdf), create_graph=True)[0]
    loss = ((normal.norm(dim=-1) - 1.0) ** 2).mean()
    loss.backward()
    g_tab = emb.hashgrid_encode_tcnn.params.grad.cpu().numpy().copy()
    g_w = lin.weight.grad.cpu().numpy().copy()
    # oracle: same computation with the numpy restatement of the double backward
    og = fields.grid_geometry(6, 10, 4, float(emb.per_level_scale))
    table = emb.hashgrid_encode_tcnn.params.detach().cpu().numpy()
    unit = ((x.detach() - emb.min_xyz) / (emb.max_xyz - emb.min_xyz)).cpu().numpy()
    w = lin.weight.detach().cpu().numpy()[0]
    y, dydx = fields.grid_encode(unit, table.reshape(-1, 2), og, 2, with_dydx=True)
    scale = 1.0 / (emb.max_xyz - emb.min_xyz).cpu().numpy()           # d unit / d x
    dy = np.tile(w[3:], (500, 1))                                      # dL/d enc for sdf.sum()
    n_o = np.einsum('nk,nkd->nd', dy.astype(np.float64), dydxo := dydx.astype(np.float64)) * scale + w[:3]
    np.testing.assert_allclose(normal.detach().cpu().numpy(), n_o, rtol=2e-3, atol=2e-3 * np.abs(n_o).max())
    norm = np.linalg.norm(n_o, axis=1)
    v = (2.0 * (norm - 1.0) / norm / 500.0)[:, None] * n_o * scale    # dL/d(d enc-part of normal / d unit)
    gto, d_dyo, _ = fields.grid_double_backward(unit, table.reshape(-1, 2), dy, v, og, 2)
    np.testing.assert_allclose(g_tab, gto.reshape(-1), rtol=5e-3, atol=5e-3 * np.abs(gto).max())
    gw_o = np.concatenate([(2.0 * (norm - 1.0) / norm / 500.0)[:, None] * n_o, np.zeros((500, 12))], 1).sum(0)
    gw_o[3:] += d_dyo.sum(0)
    np.testing.assert_allclose(g_w[0], gw_o, rtol=5e-3, atol=5e-3 * np.abs(gw_o).max())


def test_sh_and_truncexp_modules(cuda):
    emb = SHEmbedder(n_freqs=4, include_input=True, dtype='torch.float32', backend='tcnn').to(cuda)
    d = render_cases.unit_dirs(100, 2)
    out = emb(torch.from_numpy(d).to(cuda))
    assert out.shape == (100, 19)
    np.testing.assert_allclose(out[:, 3:].cpu().numpy(), fields.sh_encode(d, 4).astype(np.float16).astype(np.float32),
                               atol=1e-3)
    te = TruncExp()
    x = torch.linspace(-20, 20, 101, device=cuda, requires_grad=True)
    y = te(x)
    y.sum().backward()
    np.testing.assert_allclose(y.detach().cpu().numpy(), np.exp(x.detach().cpu().numpy()), rtol=1e-5)
    np.testing.assert_allclose(x.grad.cpu().numpy(), np.exp(np.clip(x.detach().cpu().numpy(), -15, 15)), rtol=1e-5)


@pytest.mark.parametrize('name', render_cases.DENSE_CASES)
def test_ray_marching_function_matches_reference(cuda, name):
    c = render_cases.dense_case(name)
    dev = lambda a: None if a is None else torch.from_numpy(a).to(cuda)
    sigma, rad = dev(c['sigma']).requires_grad_(True), dev(c['rad']).requires_grad_(True)
    r = ray_helper.ray_marching(sigma, rad, dev(c['zvals']), add_inf_z=c['add_inf_z'], white_bkg=c['white_bkg'],
                                bkg_color=dev(c['bkg']))
    for k in ('rgb', 'depth', 'mask', 'weights', 'trans_shift', 'alpha'):
        g = GOLD['dense_%s_%s' % (name, k)]
        assert tuple(r[k].shape) == g.shape
        np.testing.assert_allclose(r[k].detach().cpu().numpy(), g, atol=1e-5, rtol=1e-5)
    m = c['sigma'].shape[1] - (0 if c['add_inf_z'] else 1)
    assert r['sigma'].shape[1] == m and r['zvals'].shape[1] == m and r['radiance'].shape[1] == m
    ((r['rgb'] * dev(c['g_rgb'])).sum() + (r['depth'] * dev(c['g_depth'])).sum() + (r['mask'] * dev(c['g_mask'])).sum()).backward()
    gs, gr = GOLD['dense_%s_dsigma' % name], GOLD['dense_%s_drad' % name]
    np.testing.assert_allclose(sigma.grad.cpu().numpy(), gs, atol=1e-4 * np.abs(gs).max(), rtol=1e-4)
    np.testing.assert_allclose(rad.grad.cpu().numpy(), gr, atol=1e-4 * np.abs(gr).max(), rtol=1e-4)
    w = ray_helper.ray_marching(sigma.detach(), None, dev(c['zvals']), add_inf_z=c['add_inf_z'], weights_only=True)
    assert list(w.keys()) == ['weights']


def test_neus_functions(cuda):
    dev = lambda a: torch.from_numpy(a).to(cuda)
    c = render_cases.alpha_case()
    t, w = ray_helper.alpha_to_weights(dev(c['alpha']))
    np.testing.assert_allclose(t.cpu().numpy(), GOLD['alpha_a2w_trans'], atol=1e-6)
    np.testing.assert_allclose(w.cpu().numpy(), GOLD['alpha_a2w_weights'], atol=1e-6)
    alpha, rad = dev(c['alpha']).requires_grad_(True), dev(c['rad']).requires_grad_(True)
    r = ray_helper.ray_marching(dev(c['sdf']), rad, dev(c['zvals']), alpha=alpha, bkg_color=dev(c['bkg']))
    ((r['rgb'] * dev(c['g_rgb'])).sum() + (r['weights'] * dev(c['g_w'])).sum()).backward()
    np.testing.assert_allclose(alpha.grad.cpu().numpy(), GOLD['alpha_dalpha'], atol=1e-4 * np.abs(GOLD['alpha_dalpha']).max(), rtol=1e-4)
    c = render_cases.sdf_case()
    sdf, slope = dev(c['sdf']).requires_grad_(True), dev(c['slope']).requires_grad_(True)
    s = torch.tensor(float(c['s']), device=cuda, requires_grad=True)
    a = ray_helper.sdf_to_alpha(sdf, dev(c['zvals']), slope, s)
    (a * dev(c['g'])).sum().backward()
    np.testing.assert_allclose(a.detach().cpu().numpy(), GOLD['sdf_alpha'], atol=2e-6)
    np.testing.assert_allclose(sdf.grad.cpu().numpy(), GOLD['sdf_dsdf'], atol=1e-4 * np.abs(GOLD['sdf_dsdf']).max(), rtol=1e-4)
    np.testing.assert_allclose(float(s.grad), float(GOLD['sdf_ds']), rtol=1e-3)
    c = render_cases.pdf_case()
    smp = ray_helper.sample_pdf(dev(c['bins']), dev(c['weights']), c['n_sample'], det=True)
    # the cdf comes from a GPU cumsum here (different summation order than the CPU golden): 1e-4 on values in [2, 6]
    np.testing.assert_allclose(smp.cpu().numpy(), GOLD['pdf_samples'], atol=1e-4)
    rnd = ray_helper.sample_pdf(dev(c['bins']), dev(c['weights']), 64, det=False)
    assert rnd.shape == (c['bins'].shape[0], 64) and bool((rnd[:, 1:] >= rnd[:, :-1]).all())
