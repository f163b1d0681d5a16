"""CUDA depth-sample helpers and object bounds against the outputs of the reference's own
functions (tests/golden/zvals_golden.npz) and against oracle/zvals.py on seeded inputs."""
import os

import numpy as np
import pytest
import torch

from oracle import zvals as oz

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'zvals_golden.npz'))
T = lambda k: torch.from_numpy(G[k])
D = lambda k: torch.from_numpy(G[k]).cuda()


def test_sphere_ray_intersection():
    from arcnerf_b200.plugin.ray_helper import sphere_ray_intersection
    near, far, pts, mask = sphere_ray_intersection(D('rays_o'), D('rays_d'), D('sph_radius'))
    assert torch.equal(mask.cpu(), T('sph_mask'))
    assert torch.allclose(near.cpu(), T('sph_near'), atol=2e-6) and torch.allclose(far.cpu(), T('sph_far'), atol=2e-6)
    assert torch.allclose(pts.cpu(), T('sph_pts'), atol=1e-5)
    # larger seeded case against the oracle
    g = torch.Generator().manual_seed(5)
    o = torch.randn((5000, 3), generator=g) * 2
    d = torch.nn.functional.normalize(torch.randn((5000, 3), generator=g), dim=1)
    r = torch.tensor([0.5, 1.0, 3.0])
    n0, f0, p0, m0 = oz.sphere_ray_intersection(o, d, r)
    n1, f1, p1, m1 = sphere_ray_intersection(o.cuda(), d.cuda(), r.cuda())
    # grazing rays may flip the mask through the 1e-5 snapping of intermediate values: allow a handful
    assert float((m1.cpu() != m0).float().mean()) < 1e-3
    same = (m1.cpu() == m0)
    assert torch.allclose(n1.cpu()[same], n0[same], atol=1e-4) and torch.allclose(f1.cpu()[same], f0[same], atol=1e-4)


def test_near_far_from_rays():
    from arcnerf_b200.plugin.ray_helper import get_near_far_from_rays
    o, d, b = D('rays_o'), D('rays_d'), D('bounds')
    cases = dict(nf_radius=dict(bounding_radius=2.0), nf_bounds=dict(bounds=b), nf_bounds_radius=dict(bounds=b, bounding_radius=2.0),
                 nf_hard=dict(near_hardcode=2.0, far_hardcode=6.0), nf_near_hard=dict(bounding_radius=3.5, near_hardcode=0.5))
    for tag, kw in cases.items():
        near, far = get_near_far_from_rays(o, d, **kw)
        assert torch.allclose(near.cpu(), T(tag + '_near'), atol=2e-6), tag
        assert torch.allclose(far.cpu(), T(tag + '_far'), atol=2e-6), tag
    with pytest.raises(NotImplementedError):
        get_near_far_from_rays(o, d)


def test_zvals_from_near_far():
    from arcnerf_b200.plugin.ray_helper import get_zvals_from_near_far
    b = D('bounds')
    near, far = b[:, 0:1], b[:, 1:2]
    for n_pts in (64, 7):
        for inc in (True, False):
            for inv in (False, True):
                key = 'zv_%d_%d_%d' % (n_pts, inc, inv)
                z = get_zvals_from_near_far(near, far, n_pts, inclusive=inc, inverse_linear=inv)
                # the golden values come from torch's CPU linspace, whose vectorised form rounds t differently in the last
                # bit than the start + step*k form (the one torch's CUDA linspace and this kernel use): allow 2 ulp
                assert torch.allclose(z.cpu(), T(key), rtol=3e-7, atol=0), key
                z = get_zvals_from_near_far(near, far, n_pts, inclusive=inc, inverse_linear=inv, perturb=True, u=D('u_%d' % n_pts))
                assert torch.allclose(z.cpu(), T(key + '_p'), rtol=3e-7, atol=0), key


def test_fix_step():
    from arcnerf_b200.plugin.ray_helper import get_zvals_from_near_far_fix_step
    near, far = D('fix_near'), D('fix_far')
    for n_pts, fix_t in ((48, 0.05), (33, 0.011)):
        for inc in (True, False):
            z, m = get_zvals_from_near_far_fix_step(near, far, fix_t, n_pts, inclusive=inc, u=D('fix_u_%d' % n_pts))
            assert torch.equal(m.cpu(), T('fix_m_%d_%d' % (n_pts, inc)))
            assert torch.allclose(z.cpu(), T('fix_z_%d_%d' % (n_pts, inc)), rtol=3e-7, atol=0)
    with pytest.raises((AssertionError, RuntimeError)):
        get_zvals_from_near_far_fix_step(near, far, 0.0, 8)


def test_outside_sphere_points_and_sort():
    from arcnerf_b200.plugin.ray_helper import get_ray_points_by_zvals, get_zvals_outside_sphere, sort_rows, upsample_zvals
    o, d = D('rays_o'), D('rays_d')
    z, sr = get_zvals_outside_sphere(o, d, 32, 1.5)
    assert torch.allclose(sr.cpu(), T('shell_r'), rtol=1e-6) and torch.allclose(z.cpu(), T('shell_z'), atol=1e-5)
    z, sr = get_zvals_outside_sphere(o, d, 32, 1.5, perturb=True, u=D('shell_u'))
    assert torch.allclose(z.cpu(), T('shell_z_p'), atol=1e-5)
    pts = get_ray_points_by_zvals(o, d, z)
    ref = T('rays_o')[:, None, :] + T('rays_d')[:, None, :] * z.cpu()[:, :, None]
    assert torch.allclose(pts.cpu(), ref, atol=1e-6)
    for n in (1, 5, 192, 1000, 4097):
        v = torch.randn((37, n), device='cuda')
        assert torch.equal(sort_rows(v), torch.sort(v, -1)[0])
    with pytest.raises(RuntimeError, match='8192'):
        sort_rows(torch.zeros((2, 9000), device='cuda'))
    # hierarchical up-sampling: coarse samples + importance samples, sorted, importance near the weight peak
    zc = torch.linspace(2.0, 6.0, 64, device='cuda')[None].repeat(16, 1)
    w = torch.exp(-0.5 * ((zc - 4.0) / 0.1) ** 2)
    zall, new = upsample_zvals(zc, w, 128, det=True)
    assert zall.shape == (16, 192) and bool((zall[:, 1:] >= zall[:, :-1]).all())
    assert float((new - 4.0).abs().median()) < 0.2


def _density_sphere(radius=0.6, value=50.0):
    def est(dt, pts):
        return (pts.norm(dim=1) < radius).float() * value * dt
    return est


def test_volume_bound_sampling_and_optimize():
    from arcnerf_b200.plugin.obj_bound import build_obj_bound
    cfgs = dict(obj_bound=dict(volume=dict(n_grid=32, side=2.0), epoch_optim=16, epoch_optim_warmup=64, ray_sample_acc=True,
                               ray_sample_fix_step=True, near_distance=0.2, opa_thres=0.01))
    bound, kind = build_obj_bound(cfgs)
    assert kind == 'volume' and type(bound).__name__ == 'VolumeBound'
    bound = bound.cuda()
    o, d = D('rays_o'), D('rays_d')
    near, far, mask_rays = bound.get_near_far_from_rays(dict(rays_o=o, rays_d=d))
    assert near.shape == (64, 1) and mask_rays.shape == (64,) and mask_rays.dtype == torch.bool
    bound.optimize(cur_epoch=5, n_pts=128, get_est_opacity=_density_sphere())      # not a multiple of epoch_optim: no-op
    assert bool(bound.volume.get_voxel_bitfield().all())
    bound.optimize(cur_epoch=16, n_pts=128, get_est_opacity=_density_sphere())     # warm-up: every voxel
    bits = bound.volume.get_voxel_bitfield()
    centres = bound.volume.get_volume_pts().view(32, 32, 32, 3)
    inside = centres.norm(dim=-1) < 0.6 - 0.06
    outside = centres.norm(dim=-1) > 0.6 + 0.06
    assert bool(bits[inside].all()) and not bool(bits[outside].any())
    # the opacity-field rule against the reference's own result
    bound.volume.n_grid = 8
    bound.volume.register_buffer('opafield', D('opa_field0').clone())
    bound.volume.register_buffer('bitfield', torch.ones((8, 8, 8), dtype=torch.bool, device='cuda'))
    bound.volume.update_opafield_by_voxel_idx(D('opa_idx'), D('opa_new'), ema=0.95)
    bound.volume.update_bitfield_by_opafield(threshold=0.01, ops='overwrite')
    assert torch.equal(bound.volume.opafield.cpu(), T('opa_field1')) and torch.equal(bound.volume.bitfield.cpu(), T('opa_bits1'))
    bound.volume.n_grid = 32
    bound.volume.set_up_voxel_bitfield(True)
    bound.volume.set_up_voxel_opafield()
    bound.optimize(cur_epoch=16, n_pts=128, get_est_opacity=_density_sphere())
    zvals, mask_pts = bound.get_zvals_from_near_far(near, far, 128, rays_o=o, rays_d=d)
    assert zvals.shape == (64, 128) and mask_pts.dtype == torch.bool
    pts = o[:, None, :] + d[:, None, :] * zvals[:, :, None]
    assert bool((pts[mask_pts].norm(dim=1) < 0.6 + 0.12).all())
    bound.optimize(cur_epoch=80, n_pts=128, get_est_opacity=_density_sphere())     # after warm-up: partial refresh
    assert not bool(bound.volume.get_voxel_bitfield()[outside].any())


def test_bitfield_sphere_and_basic_bounds():
    from arcnerf_b200.plugin.obj_bound import build_obj_bound
    o, d = D('rays_o'), D('rays_d')
    bound, kind = build_obj_bound(dict(obj_bound=dict(bitfield=dict(n_grid=32, side=2.0), epoch_optim=16, epoch_optim_warmup=64,
                                                      near_distance=0.2)))
    assert kind == 'bitfield'
    bound = bound.cuda()
    near, far, mask_rays = bound.get_near_far_from_rays(dict(rays_o=o, rays_d=d))
    bound.optimize(16, 128, _density_sphere())
    count, ratio = bound.get_bitfield_count()
    assert 0.0 < ratio < 0.3
    zvals, mask_pts = bound.get_zvals_from_near_far(near, far, 128, rays_o=o, rays_d=d)
    pts = o[:, None, :] + d[:, None, :] * zvals[:, :, None]
    assert bool(mask_pts.any()) and bool((pts[mask_pts].norm(dim=1) < 0.6 + 0.15).all())
    bound, kind = build_obj_bound(dict(obj_bound=dict(sphere=dict(radius=1.0))))
    assert kind == 'sphere'
    near, far, mask_rays = bound.cuda().get_near_far_from_rays(dict(rays_o=o, rays_d=d))
    assert torch.equal(mask_rays.cpu(), T('sph_mask')[:, 0]) and torch.allclose(far.cpu()[:, 0], T('sph_far')[:, 0], atol=2e-6)
    bound, kind = build_obj_bound(dict())
    assert kind == 'basic'
    near, far, none = bound.get_near_far_from_rays(dict(rays_o=o, rays_d=d), bounding_radius=2.0)
    zvals, none2 = bound.get_zvals_from_near_far(near, far, 64)
    assert none is None and none2 is None and torch.allclose(near.cpu(), T('nf_radius_near'), atol=2e-6)
    assert torch.allclose(zvals[:, 0:1], near) and torch.allclose(zvals[:, -1:], far)


def test_native_volume_refresh_kernels():
    """csrc/volume_refresh.cu: candidate selection (distinct random voxels + the first occupied ones in flat order, jittered
    points inside their voxels) and the update rule against oracle/zvals.py:update_opafield (= the reference's Volume rule)."""
    import ctypes as C
    from arcnerf_b200 import _lib
    n_grid = 24
    n = n_grid ** 3
    g = torch.Generator().manual_seed(3)
    vol = (torch.rand((n, ), generator=g) < 0.2)
    vol_d = vol.cuda().contiguous()
    aabb = torch.tensor([[-1.0, -0.5, 0.0], [1.0, 1.5, 3.0]], device='cuda')
    scratch = torch.empty(_lib.load().arcb200_volume_refresh_scratch_bytes(n_grid), dtype=torch.uint8, device='cuda')
    n_uni, n_occ = n // 4, n // 4
    idx = torch.empty((n_uni + n_occ, ), dtype=torch.int32, device='cuda')
    pts = torch.empty((n_uni + n_occ, 3), dtype=torch.float32, device='cuda')
    rng = _lib.Rng()
    sets = []
    for key in (1, 2):
        _lib.call('arcb200_volume_refresh_candidates', vol_d.data_ptr(), n_grid, n_uni, n_occ, key, rng.state, rng.inc,
                  aabb.data_ptr(), scratch.data_ptr(), idx.data_ptr(), pts.data_ptr(), _lib.current_stream())
        torch.cuda.synchronize()
        uni = idx[:n_uni].cpu().long()
        assert int(uni.min()) >= 0 and int(uni.max()) < n and uni.unique().numel() == n_uni      # without replacement
        occ_ref = torch.where(vol)[0][:n_occ]
        occ = idx[n_uni:].cpu().long()
        assert torch.equal(occ[:occ_ref.numel()], occ_ref) and bool((occ[occ_ref.numel():] == -1).all())
        # every point lies inside its voxel
        valid = idx.cpu() >= 0
        v = idx.cpu().long()[valid]
        cell = torch.stack([v // (n_grid * n_grid), (v // n_grid) % n_grid, v % n_grid], 1).float()
        size = ((aabb[1] - aabb[0]) / n_grid).cpu()
        lo = aabb[0].cpu() + cell * size
        p = pts.cpu()[valid]
        assert bool((p >= lo - 1e-5).all()) and bool((p <= lo + size + 1e-5).all())
        frac = (p - lo) / size
        assert 0.45 < float(frac.mean()) < 0.55 and float(frac.std()) > 0.25                     # jitter fills the voxel
        sets.append(uni)
    assert not torch.equal(sets[0], sets[1])                                                      # key changes the subset
    # full sweep: a permutation of all voxels
    idx_all = torch.empty((n, ), dtype=torch.int32, device='cuda')
    pts_all = torch.empty((n, 3), dtype=torch.float32, device='cuda')
    _lib.call('arcb200_volume_refresh_candidates', vol_d.data_ptr(), n_grid, n, 0, 7, rng.state, rng.inc, aabb.data_ptr(),
              scratch.data_ptr(), idx_all.data_ptr(), pts_all.data_ptr(), _lib.current_stream())
    assert torch.equal(torch.sort(idx_all.cpu().long())[0], torch.arange(n))
    # update rule, duplicates included (max per voxel), frozen (negative) voxels kept
    opa0 = (torch.rand((n_grid, n_grid, n_grid), generator=g) * 0.02 - 0.002)
    cand = torch.randint(0, n, (5000, ), generator=g)
    cand[::50] = -1
    new = torch.rand((5000, ), generator=g) * 0.03
    ok = cand >= 0
    uni, inv = torch.unique(cand[ok], return_inverse=True)
    mx = torch.zeros(uni.numel()).scatter_reduce(0, inv, new[ok], 'amax', include_self=False)
    vidx = torch.stack([uni // (n_grid * n_grid), (uni // n_grid) % n_grid, uni % n_grid], 1)
    opa_ref, bits_ref = oz.update_opafield(opa0, None, vidx, mx, 0.95, 0.01)
    opa_d = opa0.cuda().contiguous()
    bits_d = torch.zeros((n, ), dtype=torch.bool, device='cuda')
    cand_d, new_d = cand.int().cuda(), new.cuda()
    _lib.call('arcb200_volume_refresh_apply', cand_d.data_ptr(), new_d.data_ptr(), 5000, 0.95, 0.01, n_grid,
              scratch.data_ptr(), opa_d.data_ptr(), bits_d.data_ptr(), _lib.current_stream())
    assert torch.equal(opa_d.cpu(), opa_ref)
    thres = min(float(opa_ref.clamp(min=0).mean()), 0.01)
    near_thres = (opa_ref.view(-1) - thres).abs() < 1e-7                 # the mean is summed in a different order
    assert torch.equal(bits_d.cpu()[~near_thres], bits_ref.view(-1)[~near_thres])


def test_volume_bound_native_and_torch_refresh_agree():
    from arcnerf_b200.plugin.obj_bound import VolumeBound
    base = dict(volume=dict(n_grid=32, side=2.0), epoch_optim=16, epoch_optim_warmup=64, ray_sample_acc=True, opa_thres=0.01)
    a = VolumeBound(dict(base)).cuda()
    b = VolumeBound(dict(base, native_refresh=False)).cuda()
    est = _density_sphere()
    for ep in (16, 32, 64, 80, 96):
        a.optimize(ep, 128, est)
        b.optimize(ep, 128, est)
    bits_a, bits_b = a.volume.get_voxel_bitfield(), b.volume.get_voxel_bitfield()
    # same field, different random points: the two occupancies differ only in the shell of voxels cut by the surface
    assert float((bits_a != bits_b).float().mean()) < 0.02
    assert abs(float(bits_a.float().mean()) - float(bits_b.float().mean())) < 0.01
