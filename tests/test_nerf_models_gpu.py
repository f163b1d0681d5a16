"""The model shells on the dense-chain networks (plugin/nerf_model.py: NeRF, NeRFPP, FullModel) against outputs of the
REFERENCE's own FullModel / NeRF / NeRFPP run on CPU with the same weights and rays (tests/golden/models_golden.npz), and
training behaviour (loss goes down through the library's backward kernels).
Tolerances: fp16 operands inside the networks -> rgb / mask abs 6e-3, depth 1e-2 of the largest depth of the batch."""
import os

import numpy as np
import pytest
import torch

import model_cases as mc
import nerf_full_cases as nc

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'models_golden.npz'))


def _load(nets, name, seed):
    nets.load_reference_state(*nc.weights(name, seed))


def _close(out, prefix, keys, suffix=''):
    for k in keys:
        ref = G['%s/%s%s' % (prefix, k, suffix)]
        got = out[k + suffix][0].detach().cpu().numpy()
        tol = 6e-3 if k != 'depth' else 1e-2 * max(1.0, float(np.abs(ref).max()))
        assert got.shape == ref.shape and np.abs(got - ref).max() < tol, (prefix, k, suffix, float(np.abs(got - ref).max()))


def _feed():
    o, d = mc.rays()
    return {'rays_o': torch.from_numpy(o).cuda()[None], 'rays_d': torch.from_numpy(d).cuda()[None]}


def test_nerf_model_matches_reference():
    from arcnerf_b200.plugin.nerf_model import FullModel, NeRF
    fg = NeRF(mc.NERF_CFG)
    _load(fg.coarse_nets, 'nerf', mc.SEEDS['fg_coarse']); _load(fg.fine_nets, 'nerf', mc.SEEDS['fg_fine'])
    model = FullModel(fg)
    with torch.no_grad():
        inf = model(_feed(), inference_only=True)
        trn = model(_feed(), inference_only=False)
    assert set(inf) == {'rgb', 'depth', 'mask'}
    assert set(trn) == {k + s for k in ('rgb', 'depth', 'mask') for s in ('_coarse', '_fine')}
    _close(inf, 'nerf/inference', ('rgb', 'depth', 'mask'))
    for s in ('_coarse', '_fine'):
        _close(trn, 'nerf/train', ('rgb', 'depth', 'mask'), s)


def test_nerfpp_full_model_matches_reference():
    from arcnerf_b200.plugin.nerf_model import FullModel, NeRF, NeRFPP
    fg, bkg = NeRF(mc.NERFPP_FG_CFG), NeRFPP(mc.NERFPP_BKG_CFG)
    _load(fg.coarse_nets, 'nerf', mc.SEEDS['fg_coarse']); _load(fg.fine_nets, 'nerf', mc.SEEDS['fg_fine'])
    _load(bkg.coarse_nets, 'nerfpp_bkg', mc.SEEDS['bkg_coarse']); _load(bkg.fine_nets, 'nerfpp_bkg', mc.SEEDS['bkg_fine'])
    model = FullModel(fg, bkg)
    with torch.no_grad():
        inf = model(_feed(), inference_only=True)
        trn = model(_feed(), inference_only=False)
    _close(inf, 'nerfpp/inference', ('rgb', 'depth', 'mask'))
    for s in ('_coarse', '_fine'):
        _close(trn, 'nerfpp/train', ('rgb', 'depth', 'mask'), s)
    # chunking over rays does not change anything
    model.fg_model.chunk_rays = 40
    with torch.no_grad():
        inf2 = model(_feed(), inference_only=True)
    assert torch.allclose(inf2['rgb'], inf['rgb'], atol=1e-6)


def test_nerf_trains_through_the_library_backward():
    """a few Adam steps on a fixed batch: the loss of both stages goes down (gradients come from arcb200_nerf_mlp_bwd and the
    compositing backward kernel)"""
    from arcnerf_b200.plugin.nerf_model import NeRF
    torch.manual_seed(0)
    cfg = dict(mc.NERF_CFG, rays=dict(mc.NERF_CFG['rays'], near=1.0, far=3.0, n_sample=32, n_importance=32))
    model = NeRF(cfg)
    o, d = mc.rays(256, seed=5)
    feed = {'rays_o': torch.from_numpy(o).cuda(), 'rays_d': torch.from_numpy(d).cuda()}
    target = torch.rand((256, 3), generator=torch.Generator().manual_seed(1)).cuda() * 0.5 + 0.25
    opt = torch.optim.Adam(model.parameters(), lr=5e-4)
    losses = []
    for _ in range(12):
        out = model(feed)
        loss = ((out['rgb_coarse'] - target) ** 2).mean() + ((out['rgb_fine'] - target) ** 2).mean()
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(float(loss.detach()))
    assert all(np.isfinite(losses)) and losses[-1] < 0.7 * losses[0], losses


def test_ngp_foreground_over_nerfpp_background_blend_and_gradients():
    """config 3 (NGP fg + NeRF++ bkg): the blend rgb_fg + T_last * rgb_bkg of full_model.py:310-330 evaluated inside the NGP
    compositing kernel equals the explicit blend of separately rendered foreground (black background) and background, and the
    background networks receive d loss / d rgb * T_last"""
    from arcnerf_b200.plugin.fg_model import NGPFgModel
    from arcnerf_b200.plugin.nerf_model import FullModel, NeRFPP
    torch.manual_seed(0)
    fg = NGPFgModel(dict(chunk_rays=512), max_rays=512, capacity=1 << 16)
    bkg = NeRFPP(mc.NERFPP_BKG_CFG)
    model = FullModel(fg, bkg)
    o, d = mc.rays(300, seed=9)
    feed = {'rays_o': torch.from_numpy(o).cuda()[None], 'rays_d': torch.from_numpy(d).cuda()[None]}
    flat = {k: v[0] for k, v in feed.items()}
    from arcnerf_b200 import _lib

    def same_jitter():                       # the march jitters its start at inference too: same generator state for every render
        fg.pipe.rng = _lib.Rng()
        for ss in fg.pipe.sets:
            ss.key = None
    with torch.no_grad():
        same_jitter()
        out = model(feed, inference_only=True)
        same_jitter()
        black = fg(dict(flat, bkg_color=torch.zeros((1, 3), device='cuda')), inference_only=True)
        b = bkg(flat, inference_only=True)
    T = fg.last_trans.clone()                # the reference's trans_shift[:, -1] (leaves the last weighted interval out)
    assert float(black['mask'].max()) > 0.05                                # the foreground does absorb something
    assert bool((T >= 1.0 - black['mask'] - 1e-4).all()) and bool((T <= 1.0 + 1e-6).all())
    assert torch.allclose(out['rgb'][0], black['rgb'] + T[:, None] * b['rgb'], atol=2e-6)
    assert torch.allclose(out['mask'][0], black['mask'], atol=1e-7)
    assert torch.allclose(out['depth'][0], black['depth'] + T * b['depth'], atol=1e-5)
    # training pass: gradients of the background nets through the fused blend == through the explicit blend
    w = torch.rand((300, 3), generator=torch.Generator().manual_seed(2)).cuda()
    model.zero_grad()
    same_jitter()
    trn = model(feed)
    (trn['rgb_coarse'][0] * w).sum().backward()
    got = {k: p.grad.clone() for k, p in bkg.named_parameters() if p.grad is not None}
    g_fg = fg.table.grad.clone()
    assert got and float(g_fg.abs().max()) > 0
    model.zero_grad()
    bt = bkg(flat)
    b_rgb = bt.get('rgb_fine', bt.get('rgb_coarse'))
    (T[:, None].detach() * b_rgb * w).sum().backward()
    for k, p in bkg.named_parameters():
        if p.grad is not None:
            ref = p.grad
            assert torch.allclose(got[k], ref, rtol=1e-3, atol=1e-4 * float(ref.abs().max()) + 1e-9), k
