"""GPU: segmented compositing, sdf_to_alpha and inverse-CDF sampling through the C ABI against the reference's
golden vectors (tests/golden/render_golden.npz) and the CPU oracle (oracle/render.py).
Tolerances: rgb/depth/mask/weights abs 1e-5 (fp32 scan order), gradients rel 1e-4 of the largest entry."""
import os

import numpy as np
import pytest
import torch

import render_cases
from arcnerf_b200 import functional as F
from oracle import render

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'render_golden.npz'))


def T(a, dev):
    return None if a is None else torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def close(a, b, atol=1e-5, rtol=1e-5):
    np.testing.assert_allclose(a.cpu().numpy() if isinstance(a, torch.Tensor) else a, b, rtol=rtol, atol=atol)


@pytest.mark.parametrize('name', render_cases.DENSE_CASES)
def test_dense_forward_backward_vs_reference_golden(cuda, name):
    c = render_cases.dense_case(name)
    sig, rad, z, bkg = T(c['sigma'], cuda), T(c['rad'], cuda), T(c['zvals'], cuda), T(c['bkg'], cuda)
    r = F.composite_forward(sig, rad, z, add_inf_z=c['add_inf_z'], white_bkg=c['white_bkg'], bkg=bkg,
                            want_weights=True, want_trans=True, want_alpha=True)
    m = c['sigma'].shape[1] - (0 if c['add_inf_z'] else 1)
    for k in ('rgb', 'depth', 'mask'):
        close(r[k], GOLD['dense_%s_%s' % (name, k)])
    for k in ('weights', 'trans_shift', 'alpha'):
        close(r[k][:, :m], GOLD['dense_%s_%s' % (name, k)])
        if m < c['sigma'].shape[1]:
            assert float(r[k][:, m:].abs().max()) == 0.0
    d_sig, d_rad = F.composite_backward(sig, rad, z, g_rgb=T(c['g_rgb'], cuda), g_depth=T(c['g_depth'], cuda),
                                        g_mask=T(c['g_mask'], cuda), add_inf_z=c['add_inf_z'],
                                        white_bkg=c['white_bkg'], bkg=bkg)
    gs, gr = GOLD['dense_%s_dsigma' % name], GOLD['dense_%s_drad' % name]
    close(d_sig, gs, atol=1e-4 * np.abs(gs).max(), rtol=1e-4)
    close(d_rad, gr, atol=1e-4 * np.abs(gr).max(), rtol=1e-4)


def test_trapezoid_fixture(cuda):
    """The reference's own synthetic density ray (tests_ray_helper.py:68-86)."""
    for append in (False, True):
        n_pts = 128
        offset = int(n_pts / 3.0); level = 20
        sigma = [0.0] * offset + [(i + 1) * level for i in range(int(offset / 5.0) + 1)]
        sigma += [(int(offset / 5.0) + 1) * level] * (3 * int((offset / 5.0) + 1))
        sigma += [(int(offset / 5.0) + 1) * level - (i + 1) * level for i in range(int(offset / 5.0) + 1)]
        sigma += [0.0] * offset
        sigma = np.array(sigma[:n_pts], np.float32)[None]
        zvals = np.linspace(2.0, 6.0, n_pts)[None]
        if append:
            zvals = np.concatenate([zvals, np.repeat(zvals[:, -1:], n_pts // 8, 1)], -1)
            sigma = np.concatenate([sigma, np.repeat(sigma[:, -1:], n_pts // 8, 1)], -1)
        zvals = zvals.astype(np.float32)
        rad = np.random.RandomState(0).uniform(0, 1, size=sigma.shape + (3,)).astype(np.float32)
        for inf in (False, True):
            for white in (False, True):
                r = F.composite_forward(T(sigma, cuda), T(rad, cuda), T(zvals, cuda), add_inf_z=inf, white_bkg=white,
                                        want_weights=True)
                key = 'trap_app%d_inf%d_white%d_' % (append, inf, white)
                m = sigma.shape[1] - (0 if inf else 1)
                close(r['rgb'], GOLD[key + 'rgb']); close(r['depth'], GOLD[key + 'depth'], atol=2e-5)
                close(r['mask'], GOLD[key + 'mask']); close(r['weights'][:, :m], GOLD[key + 'weights'])


def test_alpha_mode(cuda):
    c = render_cases.alpha_case()
    a, rad, z, bkg = T(c['alpha'], cuda), T(c['rad'], cuda), T(c['zvals'], cuda), T(c['bkg'], cuda)
    r = F.composite_forward(a, rad, z, is_alpha=True, bkg=bkg, want_weights=True, want_trans=True)
    for k in ('rgb', 'depth', 'mask', 'weights'):
        close(r[k], GOLD['alpha_%s' % k])
    close(r['trans_shift'], GOLD['alpha_a2w_trans'])
    d_a, d_rad = F.composite_backward(a, rad, z, g_rgb=T(c['g_rgb'], cuda), g_weights=T(c['g_w'], cuda), is_alpha=True,
                                      bkg=bkg)
    close(d_a, GOLD['alpha_dalpha'], atol=1e-4 * np.abs(GOLD['alpha_dalpha']).max(), rtol=1e-4)
    close(d_rad, GOLD['alpha_drad'], atol=1e-4 * np.abs(GOLD['alpha_drad']).max(), rtol=1e-4)


@pytest.mark.parametrize('seed,n_rays,max_c', [(0, 50, 40), (1, 300, 130), (2, 4096, 427), (3, 7, 1)])
def test_packed_equals_padded_dense_oracle(cuda, seed, n_rays, max_c):
    """Packed rays against the oracle run on the padded matrix FgModel would build (fg_model.py:311-316),
    including the rays that own P_max samples (T_last leaves their last interval out) and empty rays."""
    rng = np.random.RandomState(seed)
    counts = rng.randint(0, max_c + 1, n_rays).astype(np.int32)
    counts[rng.randint(0, n_rays, 3)] = max_c
    counts[0] = 0
    n = int(counts.sum())
    sig = (rng.uniform(-1, 10, n) * (rng.uniform(size=n) < 0.7)).astype(np.float32)
    rad = rng.uniform(0, 1, (n, 3)).astype(np.float32)
    t = np.concatenate([np.sort(rng.uniform(2, 6, c)) for c in counts] + [np.zeros(0)]).astype(np.float32)
    bkg = rng.uniform(0, 1, (n_rays, 3)).astype(np.float32)
    g_rgb = rng.normal(size=(n_rays, 3)).astype(np.float32)
    g_depth = rng.normal(size=n_rays).astype(np.float32)
    g_mask = rng.normal(size=n_rays).astype(np.float32)
    offsets = np.zeros(n_rays + 1, np.int32); offsets[1:] = np.cumsum(counts)
    pmax = torch.tensor([max(2, int(counts.max()))], dtype=torch.int32, device=cuda)
    r = F.composite_forward(T(sig, cuda), T(rad, cuda), T(t, cuda), offsets=T(offsets, cuda), pmax=pmax, bkg=T(bkg, cuda),
                            want_weights=True)
    S = torch.from_numpy(render.pad_packed(sig, counts)).requires_grad_(True)
    R = torch.from_numpy(render.pad_packed(rad, counts)).requires_grad_(True)
    Z = torch.from_numpy(render.pad_packed(t, counts))
    o = render.ray_marching(S, R, Z, bkg_color=torch.from_numpy(bkg))
    close(r['rgb'], o['rgb'].detach().numpy()); close(r['depth'], o['depth'].detach().numpy(), atol=2e-5)
    close(r['mask'], o['mask'].detach().numpy())
    ((o['rgb'] * torch.from_numpy(g_rgb)).sum() + (o['depth'] * torch.from_numpy(g_depth)).sum() +
     (o['mask'] * torch.from_numpy(g_mask)).sum()).backward()
    d_sig, d_rad = F.composite_backward(T(sig, cuda), T(rad, cuda), T(t, cuda), g_rgb=T(g_rgb, cuda),
                                        g_depth=T(g_depth, cuda), g_mask=T(g_mask, cuda), offsets=T(offsets, cuda),
                                        pmax=pmax, bkg=T(bkg, cuda), max_samples_per_ray=int(counts.max()))
    # oracle gradients live on the padded matrix: a packed sample collects its own entry; the last valid sample
    # also collects the padded copies of itself (they carry zero weight, hence zero gradient, except through sigma
    # of zero-length intervals, which is zero too)
    gs = np.zeros(n, np.float64); gr = np.zeros((n, 3), np.float64)
    Sg, Rg = S.grad.numpy(), R.grad.numpy()
    for i, c in enumerate(counts):
        if c:
            gs[offsets[i]:offsets[i] + c] = Sg[i, :c]; gr[offsets[i]:offsets[i] + c] = Rg[i, :c]
            gs[offsets[i] + c - 1] += Sg[i, c:].sum(); gr[offsets[i] + c - 1] += Rg[i, c:].sum(0)
    close(d_sig, gs, atol=1e-4 * max(np.abs(gs).max(), 1e-12), rtol=1e-4)
    close(d_rad, gr, atol=1e-4 * max(np.abs(gr).max(), 1e-12), rtol=1e-4)


def test_composite_limits(cuda):
    z = torch.zeros((1, 5000), device=cuda)
    with pytest.raises(RuntimeError):
        F.composite_backward(z, None, z, g_depth=torch.zeros(1, device=cuda))


def test_sdf_to_alpha_and_sample_cdf(cuda):
    c = render_cases.sdf_case()
    sdf, z, slope, g = T(c['sdf'], cuda), T(c['zvals'], cuda), T(c['slope'], cuda), T(c['g'], cuda)
    s = torch.tensor([float(c['s'])], device=cuda)
    for sv in (s, float(c['s'])):
        a = F.sdf_to_alpha(sdf, z, slope, sv)
        close(a, GOLD['sdf_alpha'], atol=2e-6)
    d_sdf, d_slope, d_s = F.sdf_to_alpha_backward(sdf, z, slope, s, g)
    close(d_sdf, GOLD['sdf_dsdf'], atol=1e-4 * np.abs(GOLD['sdf_dsdf']).max(), rtol=1e-4)
    close(d_slope, GOLD['sdf_dslope'], atol=1e-4 * np.abs(GOLD['sdf_dslope']).max(), rtol=1e-4)
    close(d_s, GOLD['sdf_ds'].reshape(1), atol=1e-4 * abs(float(GOLD['sdf_ds'])), rtol=1e-3)
    c = render_cases.pdf_case()
    bins, w = T(c['bins'], cuda), T(c['weights'], cuda)
    w = w + 1e-5
    pdf = w / w.sum(-1, keepdim=True)
    cdf = torch.cat([torch.zeros_like(pdf[:, :1]), torch.cumsum(pdf, -1)], -1).contiguous()
    u = torch.linspace(0.0, 1.0, steps=c['n_sample'], device=cuda)
    got = torch.sort(F.sample_cdf_gather(bins, cdf, u), -1)[0]
    close(got, GOLD['pdf_samples'], atol=2e-5)
    u2 = u.expand(bins.shape[0], -1).contiguous()
    assert torch.equal(F.sample_cdf_gather(bins, cdf, u2), F.sample_cdf_gather(bins, cdf, u))


@pytest.mark.parametrize('mode', ['packed_bkg', 'packed_white', 'dense'])
def test_fused_composite_huber_train_equals_three_kernels(mode):
    """arcb200_composite_huber_train == composite_fwd -> huber_loss -> composite_bwd, bit for bit (loss: atomics order)"""
    from arcnerf_b200 import _lib
    g = torch.Generator().manual_seed(17)
    R = 777
    if mode == 'dense':
        P = 67
        counts = torch.full((R, ), P, dtype=torch.int32)
    else:
        counts = torch.randint(0, 200, (R, ), generator=g, dtype=torch.int32)
        counts[::13] = 0
        counts[5] = 1
        counts[7] = 199
    offsets = torch.zeros(R + 1, dtype=torch.int32)
    offsets[1:] = torch.cumsum(counts, 0)
    N = int(offsets[-1])
    t = torch.cat([torch.sort(2.0 + 4.0 * torch.rand(int(c), generator=g))[0] for c in counts]) if N else torch.zeros(0)
    sigma = (torch.rand(N, generator=g) * 30.0 * (torch.rand(N, generator=g) > 0.4)).float()
    rad = torch.rand((N, 3), generator=g)
    target = torch.rand((R, 3), generator=g)
    bkg = torch.rand((R, 3), generator=g) if mode == 'packed_bkg' else None
    white = int(mode == 'packed_white')
    dev = 'cuda'
    off_d = offsets.to(dev) if mode != 'dense' else None
    pmax = torch.tensor([int(counts.max())], dtype=torch.int32, device=dev)
    sig_d, rad_d, t_d, tgt_d = sigma.to(dev), rad.to(dev), t.to(dev), target.to(dev)
    bkg_d = bkg.to(dev) if bkg is not None else None
    st = _lib.current_stream()
    dense_p = P if mode == 'dense' else 0
    rgb1, depth, acc = (torch.empty((R, 3), device=dev), torch.empty(R, device=dev), torch.empty(R, device=dev))
    _lib.call('arcb200_composite_fwd', _lib.ptr(off_d), dense_p, N, pmax.data_ptr(), sig_d.data_ptr(), 0, rad_d.data_ptr(),
              t_d.data_ptr(), None, 0, white, _lib.ptr(bkg_d), int(bkg is not None), rgb1.data_ptr(), depth.data_ptr(), acc.data_ptr(),
              None, None, None, R, st)
    loss1, g_rgb1 = torch.zeros(1, device=dev), torch.empty((R, 3), device=dev)
    _lib.call('arcb200_huber_loss', rgb1.data_ptr(), tgt_d.data_ptr(), 0.1, 3000.0, g_rgb1.data_ptr(), loss1.data_ptr(), R, st)
    dg1, dr1 = torch.full((N, ), 7.0, device=dev), torch.full((N, 3), 7.0, device=dev)
    _lib.call('arcb200_composite_bwd', _lib.ptr(off_d), dense_p, N, pmax.data_ptr(), sig_d.data_ptr(), 0, rad_d.data_ptr(),
              t_d.data_ptr(), None, 0, white, _lib.ptr(bkg_d), int(bkg is not None), g_rgb1.data_ptr(), None, None, None,
              dg1.data_ptr(), dr1.data_ptr(), 256, R, st)
    rgb2, g_rgb2, loss2 = torch.empty((R, 3), device=dev), torch.empty((R, 3), device=dev), torch.zeros(1, device=dev)
    dg2, dr2 = torch.full((N, ), 9.0, device=dev), torch.full((N, 3), 9.0, device=dev)
    counter = torch.tensor([41], dtype=torch.int32, device=dev)
    _lib.call('arcb200_composite_huber_train', _lib.ptr(off_d), dense_p, N, pmax.data_ptr(), sig_d.data_ptr(), rad_d.data_ptr(),
              t_d.data_ptr(), None, 0, white, _lib.ptr(bkg_d), int(bkg is not None), tgt_d.data_ptr(), 0.1, 3000.0, rgb2.data_ptr(),
              g_rgb2.data_ptr(), loss2.data_ptr(), dg2.data_ptr(), dr2.data_ptr(), counter.data_ptr(), 256, R, st)
    torch.cuda.synchronize()
    assert torch.equal(rgb1, rgb2) and torch.equal(g_rgb1, g_rgb2)
    assert torch.equal(dg1, dg2) and torch.equal(dr1, dr2)
    assert abs(float(loss1) - float(loss2)) <= 2e-6 * abs(float(loss1))
    assert float(dg2.abs().max()) > 0 and int(counter) == 42


def test_ray_marching_reports_decreasing_zvals_without_stalling():
    """the reference asserts `zvals is not all increase....` (ray_helper.py:532-534) with a host synchronisation; here the check is a
    device flag read one call later (or by check_zvals_assert), immediately under ARCB200_STRICT_ASSERTS=1"""
    from arcnerf_b200.plugin import ray_helper as rh
    good = torch.linspace(1.0, 2.0, 16, device='cuda')[None].repeat(8, 1).contiguous()
    bad = good.clone()
    bad[3, 7] = 0.5
    sigma, rad = torch.rand((8, 16), device='cuda'), torch.rand((8, 16, 3), device='cuda')
    rh.check_zvals_assert()
    rh.ray_marching(sigma, rad, good)
    rh.check_zvals_assert()                                  # nothing pending fails
    rh.ray_marching(sigma, rad, bad)                         # queued, not raised yet
    with pytest.raises(AssertionError, match='zvals is not all increase'):
        rh.check_zvals_assert()
    rh.ray_marching(sigma, rad, good)
    rh.check_zvals_assert()                                  # the flag was cleared by the raise
    nearly = good.clone()
    nearly[2, 5] = nearly[2, 4] - 5e-6                       # |delta| < 1e-5 counts as zero (ray_helper.py:533)
    rh.ray_marching(sigma, rad, nearly)
    rh.check_zvals_assert()
