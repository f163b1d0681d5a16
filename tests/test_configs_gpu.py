"""Parity cases for the other BASELINE.json configurations, at the level this library covers them: the sampling and
integration chains that the reference's NeRF / NeuS / NeRF++ foreground and background models run around their networks
(which are stood in for by analytic fields, like the reference's own FakeGeoNet fixture, tests_fg_model.py:293-309).
CUDA path (plugin mirrors -> C ABI) against oracle/ (torch restatement pinned to the reference by tests/golden/).

  configs[0] vanilla NeRF : 64 stratified coarse samples -> weights -> 128 inverse-CDF samples, merged -> composite
  configs[2] NeuS         : SDF + slope at interval mid points -> sdf_to_alpha -> composite from alpha, gradient of the scale
  configs[3] NGP + NeRF++ : inverse-sphere shell samples for the background, blend rgb_fg + T_last * rgb_bkg
  configs[4] 1920x1080 frame + 512^3 density grid: tests/test_configs_gpu.py::test_config4_*
"""
import numpy as np
import pytest
import torch

from oracle import render as orc
from oracle import zvals as oz

pytestmark = pytest.mark.gpu


def _rays(n, seed):
    from arcnerf_b200 import synth
    o, d = synth.train_ray_batch(n, seed=seed)
    return torch.from_numpy(o), torch.from_numpy(d)


def _blob_field(pts, dirs):
    """analytic density (two blobs) and view-dependent colour, fp32 torch ops only (runs on both devices)"""
    c1 = torch.tensor([0.2, -0.1, 0.1], device=pts.device)
    c2 = torch.tensor([-0.3, 0.3, -0.2], device=pts.device)
    sigma = 40.0 * torch.exp(-((pts - c1) ** 2).sum(-1) / 0.05) + 25.0 * torch.exp(-((pts - c2) ** 2).sum(-1) / 0.02)
    rgb = torch.sigmoid(torch.stack([3.0 * pts[..., 0] + dirs[..., 2], 2.0 * pts[..., 1] - dirs[..., 0],
                                     pts[..., 2] + dirs[..., 1]], -1))
    return sigma, rgb


def test_config0_vanilla_nerf_hierarchical_sampling_chain():
    from arcnerf_b200.plugin import ray_helper as rh
    n, n_coarse, n_fine = 1024, 64, 128
    o, d = _rays(n, 21)
    near, far = torch.full((n, 1), 2.0), torch.full((n, 1), 6.0)          # the Blender near/far of the reference config
    u = torch.rand((n, n_coarse), generator=torch.Generator().manual_seed(1))
    # ---- oracle chain (CPU)
    z0 = oz.get_zvals_from_near_far(near, far, n_coarse, u=u)
    p0 = o[:, None, :] + d[:, None, :] * z0[..., None]
    s0, c0 = _blob_field(p0, d[:, None, :].expand_as(p0))
    out0 = orc.ray_marching(s0, c0, z0)
    mids = 0.5 * (z0[:, 1:] + z0[:, :-1])
    uu = torch.linspace(0.0, 1.0, n_fine)[None].expand(n, n_fine).contiguous()
    znew = torch.sort(orc.sample_pdf(mids, out0['weights'][:, 1:n_coarse - 1], uu), -1)[0]
    z1 = torch.sort(torch.cat([z0, znew], -1), -1)[0]
    p1 = o[:, None, :] + d[:, None, :] * z1[..., None]
    s1, c1 = _blob_field(p1, d[:, None, :].expand_as(p1))
    out1 = orc.ray_marching(s1, c1, z1)
    # ---- CUDA chain
    od, dd = o.cuda(), d.cuda()
    g0 = rh.get_zvals_from_near_far(near.cuda(), far.cuda(), n_coarse, perturb=True, u=u.cuda())
    assert torch.allclose(g0.cpu(), z0, rtol=3e-7, atol=0)
    gp0 = rh.get_ray_points_by_zvals(od, dd, g0)
    gs0, gc0 = _blob_field(gp0, dd[:, None, :].expand_as(gp0))
    gout0 = rh.ray_marching(gs0, gc0, g0)
    assert torch.allclose(gout0['weights'].cpu(), out0['weights'], atol=2e-6)
    g1, gnew = rh.upsample_zvals(g0, gout0['weights'], n_fine, det=True)
    assert g1.shape == (n, n_coarse + n_fine) and bool((g1[:, 1:] >= g1[:, :-1]).all())
    # inverse CDF: where the coarse weights vanish the CDF is nearly flat, so 1e-7 of weight noise moves a sample by 1e-4
    dz = (gnew.cpu() - znew).abs()
    assert float(dz.median()) < 1e-5 and float(dz.quantile(0.999)) < 5e-3 and float(dz.max()) < 0.1
    gp1 = rh.get_ray_points_by_zvals(od, dd, g1)
    gs1, gc1 = _blob_field(gp1, dd[:, None, :].expand_as(gp1))
    gout1 = rh.ray_marching(gs1, gc1, g1)
    for k, tol in (('rgb', 2e-4), ('depth', 1e-3), ('mask', 2e-4)):
        assert torch.allclose(gout1[k].cpu(), out1[k], atol=tol), k
    assert float(out1['mask'].max()) > 0.5                                 # the rays do hit the blobs


def test_config2_neus_sdf_to_alpha_chain_and_scale_gradient():
    from arcnerf_b200.plugin import ray_helper as rh
    n, p = 512, 96
    o, d = _rays(n, 22)
    near, far, _, hit = oz.sphere_ray_intersection(o, d, torch.tensor([1.2]))
    far = torch.where(far <= near, near + 1.0, far)
    z = oz.get_zvals_from_near_far(near, far, p + 1)
    mid = 0.5 * (z[:, 1:] + z[:, :-1])
    pts = o[:, None, :] + d[:, None, :] * mid[..., None]
    radius = 0.8
    sdf = pts.norm(dim=-1) - radius
    normal = pts / pts.norm(dim=-1, keepdim=True)
    slope = (normal * d[:, None, :]).sum(-1)
    col = torch.sigmoid(2.0 * normal)
    s_ref = torch.tensor(64.0, requires_grad=True)
    a_ref = orc.sdf_to_alpha(sdf, z, slope, s_ref)
    out_ref = orc.ray_marching(None, col, mid, alpha=a_ref)
    (out_ref['rgb'].sum() + out_ref['depth'].sum()).backward()
    s_gpu = torch.tensor(64.0, device='cuda', requires_grad=True)
    a_gpu = rh.sdf_to_alpha(sdf.cuda(), z.cuda(), slope.cuda(), s_gpu)
    out_gpu = rh.ray_marching(None, col.cuda(), mid.cuda(), alpha=a_gpu)
    (out_gpu['rgb'].sum() + out_gpu['depth'].sum()).backward()
    assert torch.allclose(a_gpu.detach().cpu(), a_ref.detach(), atol=2e-6)
    for k in ('rgb', 'depth', 'mask', 'weights', 'trans_shift'):
        assert torch.allclose(out_gpu[k].detach().cpu(), out_ref[k].detach(), atol=3e-5), k
    assert abs(float(s_gpu.grad) - float(s_ref.grad)) <= 2e-3 * abs(float(s_ref.grad)) + 1e-6
    hit_rays = hit[:, 0] & (out_ref['mask'].detach() > 0.5)
    assert int(hit_rays.sum()) > 50
    # rendered depth of rays that hit the sphere = distance to the surface (what NeuS's alpha is built to give)
    n_s, _, _, m_s = oz.sphere_ray_intersection(o, d, torch.tensor([radius]))
    sel = hit_rays & m_s[:, 0]
    assert float((out_gpu['depth'].detach().cpu()[sel] - n_s[sel, 0]).abs().median()) < 0.05


def test_config3_nerfpp_background_shells_and_blend():
    from arcnerf_b200.plugin import ray_helper as rh
    n, n_bkg = 768, 32
    o, d = _rays(n, 23)
    # foreground inside the unit sphere (stand-in for the NGP foreground), background on inverse-sphere shells
    near, far = oz.get_near_far_from_rays(o, d, bounding_radius=1.0)
    zf = oz.get_zvals_from_near_far(near, far, 48)
    pf = o[:, None, :] + d[:, None, :] * zf[..., None]
    sf, cf = _blob_field(pf, d[:, None, :].expand_as(pf))
    fg = orc.ray_marching(sf, cf, zf)
    ub = torch.rand((1, n_bkg), generator=torch.Generator().manual_seed(3))
    zb, radii = oz.get_zvals_outside_sphere(o, d, n_bkg, 1.0, u=ub)
    pb = o[:, None, :] + d[:, None, :] * zb[..., None]
    pb4 = torch.cat([pb / radii[None, :, None], (1.0 / radii)[None, :, None].expand(n, n_bkg, 1)], -1)   # (x/r, 1/r)
    sb = 2.0 + torch.sin(3.0 * pb4[..., 0]) + pb4[..., 3]
    cb = torch.sigmoid(pb4[..., :3] * 2.0)
    bk = orc.ray_marching(sb, cb, zb, add_inf_z=True)
    rgb_ref = fg['rgb'] + fg['trans_shift'][:, -1:] * bk['rgb']
    depth_ref = fg['depth'] + fg['trans_shift'][:, -1] * bk['depth']
    od, dd = o.cuda(), d.cuda()
    gnear, gfar = rh.get_near_far_from_rays(od, dd, bounding_radius=1.0)
    assert torch.allclose(gnear.cpu(), near, atol=2e-6) and torch.allclose(gfar.cpu(), far, atol=2e-6)
    gzf = rh.get_zvals_from_near_far(gnear, gfar, 48)
    gpf = rh.get_ray_points_by_zvals(od, dd, gzf)
    gsf, gcf = _blob_field(gpf, dd[:, None, :].expand_as(gpf))
    gfg = rh.ray_marching(gsf, gcf, gzf)
    gzb, gradii = rh.get_zvals_outside_sphere(od, dd, n_bkg, 1.0, perturb=True, u=ub.cuda())
    assert torch.allclose(gradii.cpu(), radii, rtol=1e-6) and torch.allclose(gzb.cpu(), zb, atol=2e-5)
    gpb = rh.get_ray_points_by_zvals(od, dd, gzb)
    gpb4 = torch.cat([gpb / gradii[None, :, None], (1.0 / gradii)[None, :, None].expand(n, n_bkg, 1)], -1)
    gsb = 2.0 + torch.sin(3.0 * gpb4[..., 0]) + gpb4[..., 3]
    gcb = torch.sigmoid(gpb4[..., :3] * 2.0)
    gbk = rh.ray_marching(gsb, gcb, gzb, add_inf_z=True)
    rgb = gfg['rgb'] + gfg['trans_shift'][:, -1:] * gbk['rgb']            # FullModel.blend_bkg_rgb (full_model.py:310-330)
    depth = gfg['depth'] + gfg['trans_shift'][:, -1] * gbk['depth']
    assert torch.allclose(rgb.cpu(), rgb_ref, atol=1e-4)
    assert torch.allclose(depth.cpu(), depth_ref, rtol=1e-4, atol=1e-3)
    assert float((fg['trans_shift'][:, -1] > 0.5).float().mean()) > 0.1    # the background does show through


def _pipe(max_rays, capacity):
    from arcnerf_b200 import synth
    from arcnerf_b200.ngp import NGP_DEFAULTS, NGPPipeline, init_params
    params = init_params(dict(NGP_DEFAULTS), seed=11, grid_scale=0.5)
    pipe = NGPPipeline(max_rays=max_rays, capacity=capacity, params=params)
    pipe.set_occupancy(synth.occupancy_volume('lego'))
    return pipe, params


def test_config4_density_grid_tiles_and_oracle():
    """512^3 density grid for mesh extraction, split by voxel slab: slabs concatenate to the full grid (what two ranks
    would gather), the 134 M-point grid runs at full size, and sampled voxels agree with the CPU oracle's field."""
    from arcnerf_b200.dist import shard_range
    from arcnerf_b200.ngp import NGP_DEFAULTS
    from oracle.ngp import NGPOracle
    pipe, params = _pipe(4096, 1 << 19)
    full = pipe.density_grid(64)
    parts = [pipe.density_grid(64, x_range=shard_range(64, r, 3)) for r in range(3)]
    assert torch.equal(torch.cat(parts, 0), full) and full.shape == (64, 64, 64)
    big = pipe.density_grid(512, slab=32)
    assert big.shape == (512, 512, 512) and bool(torch.isfinite(big).all()) and float(big.min()) > 0.0
    # voxel (i,j,k) of the 512^3 grid against the oracle at its centre
    g = torch.Generator().manual_seed(4)
    ijk = torch.randint(0, 512, (1500, 3), generator=g)
    unit = ((ijk.float() + 0.5) / 512.0).numpy().astype(np.float32)
    orcl = NGPOracle({k: v.numpy() for k, v in params.items()}, dict(NGP_DEFAULTS))
    sig, _, _ = orcl.field(unit, np.tile(np.array([[0.0, 0.0, 1.0]], np.float32), (1500, 1)))
    got = big[ijk[:, 0], ijk[:, 1], ijk[:, 2]].cpu()
    rel = (got - sig).abs() / sig.abs().clamp_min(1e-3)
    assert float(rel.median()) < 5e-3 and float((rel > 5e-2).float().mean()) < 0.01    # fp16 features through exp()


def test_config4_full_hd_frame_in_chunks():
    """1920x1080 frame (2 073 600 rays) through the chunked renderer; a row band rendered on its own (a rank's shard)
    agrees with the same rows of the full frame up to the march's per-launch start jitter."""
    from arcnerf_b200 import synth
    from arcnerf_b200.dist import shard_range
    pipe, _ = _pipe(32768, 3 << 19)
    o, d = synth.camera_rays(1920, 1080, synth.sphere_poses(1, seed=2)[0])
    o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    white = torch.ones((1, 3), device='cuda')
    rgb, depth, mask = pipe.render(o, d, white)
    assert rgb.shape == (1920 * 1080, 3) and bool(torch.isfinite(rgb).all())
    assert 0.02 < float((mask > 0.5).float().mean()) < 0.9
    a, b = shard_range(o.shape[0], 1, 8)
    r1, d1, m1 = pipe.render(o[a:b], d[a:b], white)
    assert float((r1 - rgb[a:b]).abs().mean()) < 0.02
    assert float(((m1 == 0) == (mask[a:b] == 0)).float().mean()) > 0.995      # grazing rays: the jitter decides
