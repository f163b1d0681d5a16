"""MultiVol background mirror (plugin/multivol_bkg_model.py; reference arcnerf/models/multivol_bkg_model.py:25-261):
cascaded sampling outside the inner volume, forward through stand-in networks against the oracle's ray_marching on the
same samples, and the density-grid refresh."""
import pytest
import torch

from oracle import render as orc

pytestmark = pytest.mark.gpu


class _Geo(torch.nn.Module):
    """analytic density: a shell of matter between radius 2 and 3 (outside the inner unit volume)"""

    def forward(self, pts):
        r = pts.norm(dim=-1, keepdim=True)
        return 8.0 * ((r > 2.0) & (r < 3.0)).float(), torch.sigmoid(pts)


class _Rad(torch.nn.Module):
    def forward(self, x, view_dirs, normals, feat):
        return torch.sigmoid(feat * 2.0 + view_dirs)


def _model(**kw):
    from arcnerf_b200.plugin.multivol_bkg_model import MultiVol
    return MultiVol(_Geo(), _Rad(), dict(n_grid=32, side=2.0, origin=(0.0, 0.0, 0.0), n_cascade=4), n_sample=256,
                    cone_angle=0.004, epoch_optim=16, epoch_optim_warmup=64, **kw).cuda()


def _rays(n):
    g = torch.Generator().manual_seed(2)
    o = (torch.rand((n, 3), generator=g) - 0.5) * 1.0
    d = torch.nn.functional.normalize(torch.randn((n, 3), generator=g), dim=1)
    return o.cuda(), d.cuda()


def test_multivol_sampling_forward_and_refresh():
    m = _model()
    assert m.total_n_elements == 32 ** 3 * 3 and m.max_volume.get_len() == (16.0, 16.0, 16.0)
    o, d = _rays(600)
    near, far = m.get_near_far_from_rays(o, d)
    assert near.shape == (600, 1) and bool((near == 0).all()) and bool((far > 7.0).all()) and bool((far < 14.0).all())
    exit_pts = o + d * far
    assert torch.allclose(exit_pts.abs().amax(dim=1), torch.full((600,), 8.0, device='cuda'), atol=1e-4)
    zvals, mask = m.get_zvals_from_near_far(near, far, 256, o, d)
    pts = o[:, None, :] + d[:, None, :] * zvals[..., None]
    inside_inner = pts[mask].abs().amax(dim=1) < 1.0 - 1e-3
    assert bool(mask.any()) and not bool(inside_inner.any())          # the inner volume belongs to the foreground model
    out = m({'rays_o': o, 'rays_d': d})
    assert set(out) == {'rgb', 'depth', 'mask'} and out['rgb'].shape == (600, 3)
    # same samples through the oracle's compositing
    torch.manual_seed(0)
    m2 = _model()
    from arcnerf_b200.plugin import multivol_func
    multivol_func.reset_rng()
    z2, k2 = m2.get_zvals_from_near_far(near, far, 256, o, d)
    p = max(1, int(k2.sum(dim=1).max()))
    z2, k2 = z2[:, :p], k2[:, :p]
    sig, rad = m2.get_sigma_radiance_by_mask_pts(o, d, z2, k2)
    ref = orc.ray_marching(sig.cpu(), rad.cpu(), z2.cpu())
    multivol_func.reset_rng()
    out2 = m2({'rays_o': o, 'rays_d': d})
    for k in ('rgb', 'depth', 'mask'):
        assert torch.allclose(out2[k].cpu(), ref[k], atol=2e-5), k
    assert float(out2['mask'].mean()) > 0.5                              # the shell is hit
    # refresh: after the warm-up sweep the cells far from the shell are pruned, sampling concentrates on it
    before = int(mask.sum())
    m.optimize(5)
    assert m.ema_step == 0
    m.optimize(16)
    assert m.ema_step == 1 and 0 < m.get_density_grid_mean()
    zv, mk = m.get_zvals_from_near_far(near, far, 256, o, d)
    pts = (o[:, None, :] + d[:, None, :] * zv[..., None])[mk]
    r = pts.norm(dim=1)
    assert int(mk.sum()) < before and float(((r > 1.4) & (r < 3.9)).float().mean()) > 0.9
    m.optimize(80)                                                        # steady-state form (n/4 + n/4 candidates)
    assert m.ema_step == 2
