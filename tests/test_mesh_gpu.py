"""Mesh extraction: csrc/mcubes.cu against the numpy restatement oracle/marching_cubes.py bit for bit, and
against closed-form surfaces.  skimage (the reference's marching cubes) is not in this image: see the oracle's header."""
import os

import numpy as np
import pytest
import torch

from oracle import marching_cubes as M
from mesh_cases import blobs_volume, closed_and_oriented, sphere_volume

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'volume_golden.npz'))


def _same(a, b):
    return a.shape == b.shape and np.array_equal(a.view(np.uint32) if a.dtype == np.float32 else a,
                                                 b.view(np.uint32) if b.dtype == np.float32 else b)


@pytest.mark.parametrize('grad_dir', ['descent', 'ascent'])
@pytest.mark.parametrize('shape', [(40, 40, 40), (33, 21, 50), (2, 2, 2), (5, 2, 3)])
def test_marching_cubes_equals_restatement(shape, grad_dir):
    from arcnerf_b200.plugin.mesh import extract_mesh
    vol, vs, vl, level = blobs_volume(shape, seed=sum(shape))
    if grad_dir == 'ascent':
        vol, level = -vol, -level
    v, f, n = extract_mesh(vol, level, vs, vl, grad_dir)
    rv, rf, rn = M.extract_mesh(vol, level, vs, vl, grad_dir)
    assert _same(v, rv) and _same(f, rf)
    assert n.shape == rn.shape and np.allclose(n, rn, atol=2e-6)            # fp32 sqrt / division sequence: last-bit agreement
    if min(shape) > 2:
        assert len(f) > 0


def test_all_256_cases_give_a_closed_oriented_surface():
    """white noise hits every sign configuration; padded with outside values the surface must be closed, every directed edge
    matched by its reverse exactly once"""
    from arcnerf_b200.plugin.mesh import extract_mesh
    rng = np.random.default_rng(3)
    vol = np.zeros((36, 36, 36), np.float32)
    vol[2:-2, 2:-2, 2:-2] = rng.random((32, 32, 32), dtype=np.float32)
    v, f, _ = extract_mesh(vol, 0.5, (1., 1., 1.), (36., 36., 36.))
    rv, rf, _ = M.extract_mesh(vol, 0.5, (1., 1., 1.), (36., 36., 36.))
    assert _same(v, rv) and _same(f, rf)
    closed_and_oriented(v, f, manifold=False)


def test_sphere_area_volume_normals():
    from arcnerf_b200.plugin.mesh import extract_mesh
    n, r = 96, 0.6
    vol, vs, vl = sphere_volume(n, r)
    v, f, nr = extract_mesh(torch.from_numpy(vol).cuda(), 50.0, vs, vl)
    assert v.is_cuda and f.dtype == torch.int32
    v, f, nr = v.cpu().numpy(), f.cpu().numpy(), nr.cpu().numpy()
    closed_and_oriented(v, f)
    d = np.linalg.norm(v, axis=1)
    assert abs(d - r).max() < 2e-3
    a, b, c = v[f[:, 0]], v[f[:, 1]], v[f[:, 2]]
    cr = np.cross(b - a, c - a)
    assert abs(0.5 * np.linalg.norm(cr, axis=1).sum() / (4 * np.pi * r * r) - 1) < 2e-3
    assert abs((a * cr).sum() / 6 / (4 / 3 * np.pi * r ** 3) - 1) < 2e-3    # positive: faces wind outward
    assert ((nr * (v / d[:, None])).sum(1)).min() > 0.999
    assert len(v) - 3 * len(f) // 2 + len(f) == 2                           # Euler characteristic of a sphere


def test_empty_and_full_volumes_and_bad_arguments():
    from arcnerf_b200.plugin.mesh import extract_mesh
    for fill in (0.0, 100.0):
        v, f, n = extract_mesh(np.full((8, 8, 8), fill, np.float32), 50.0, (1, 1, 1), (8, 8, 8))
        assert v.shape == (0, 3) and f.shape == (0, 3) and n.shape == (0, 3)
    with pytest.raises(AssertionError):
        extract_mesh(np.zeros((4, 4, 4), np.float32), 0.5, (1, 1, 1), (4, 4, 4), 'sideways')
    with pytest.raises(RuntimeError):
        extract_mesh(torch.zeros((4, 4, 4)), 0.5, (1, 1, 1), (4, 4, 4))
    with pytest.raises(RuntimeError):
        extract_mesh(np.zeros((1, 4, 4), np.float32), 0.5, (1, 1, 1), (1, 4, 4))


@pytest.mark.parametrize('tag', ['cube16', 'box9', 'one', 'box67'])
def test_volume_pts_match_reference_volume(tag):
    """the reference's Volume.get_volume_pts() (tests/golden/make_golden_volume.py)"""
    from arcnerf_b200.plugin.mesh import volume_pts
    lo, hi, n = GOLD['%s_lo' % tag], GOLD['%s_hi' % tag], int(GOLD['%s_n' % tag])
    ref = GOLD['%s_pts' % tag]
    got = volume_pts(lo, hi, n).cpu().numpy()
    assert _same(got, ref)
    if n >= 4:
        part = volume_pts(lo, hi, n, 1, 3).cpu().numpy()
        assert _same(part, got.reshape(n, n * n, 3)[1:3].reshape(-1, 3))


def test_ngp_density_grid_and_mesh_extraction():
    """run_infer_volume on the NGP pipeline (infer_func.py:448-500): the lattice density equals forward_pts of the reference's
    voxel centres, slab ranges tile the grid, and the extracted mesh is in world coordinates"""
    from arcnerf_b200.ngp import NGPPipeline
    from arcnerf_b200.plugin.mesh import volume_pts
    pipe = NGPPipeline(max_rays=256, capacity=1 << 16, grid_scale=1.0)
    n = 24
    lo, hi = pipe.aabb[0].tolist(), pipe.aabb[1].tolist()
    grid = pipe.density_grid(n, slab=5)
    ref = pipe.forward_pts(volume_pts(lo, hi, n)).view(n, n, n)
    assert torch.equal(grid, ref)
    part = pipe.density_grid(n, x_range=(7, 19), slab=4)
    assert torch.equal(part, grid[7:19])
    level = float(grid.median())
    v, f, nr = pipe.extract_mesh(n, level=level)
    assert v.shape[1] == 3 and f.shape[1] == 3 and len(f) > 0 and nr.shape == v.shape
    assert float(v.min()) >= lo[0] - 1e-5 and float(v.max()) <= hi[0] + 1e-5
    rv, rf, _ = M.extract_mesh(grid.cpu().numpy(), level, [(hi[a] - lo[a]) / n for a in range(3)], [hi[a] - lo[a] for a in range(3)])
    assert np.array_equal(f.cpu().numpy(), rf) and np.allclose(v.cpu().numpy(), rv + (np.asarray(lo) + np.asarray(hi)) / 2, atol=1e-6)


def test_mesh_components_and_colours():
    """get_mesh_components (infer_func.py:582-632): face normals / centres / mean vertex normals against numpy, and the colour
    lookup forward_pts_dir against the oracle's field (fp16 tensor-core path: the tolerance of the field tests)"""
    from arcnerf_b200.ngp import NGP_DEFAULTS, NGPPipeline, init_params
    from arcnerf_b200.plugin.mesh import extract_mesh, get_mesh_components
    from oracle.ngp import NGP_CONFIG, NGPOracle
    vol, vs, vl = sphere_volume(32, 0.6)
    v, f, _ = extract_mesh(torch.from_numpy(vol).cuda(), 50.0, vs, vl)
    vn, fn, fc = get_mesh_components(v, f)
    vh, fh = v.cpu().numpy().astype(np.float64), f.cpu().numpy()
    a, b, c = vh[fh[:, 0]], vh[fh[:, 1]], vh[fh[:, 2]]
    n = np.cross(b - a, c - a)
    n /= np.linalg.norm(n, axis=1, keepdims=True)
    assert np.allclose(fn.cpu().numpy(), n, atol=2e-4) and np.allclose(fc.cpu().numpy(), (a + b + c) / 3, atol=1e-6)
    acc = np.zeros_like(vh)
    for k in range(3):
        np.add.at(acc, fh[:, k], n)
    acc /= np.linalg.norm(acc, axis=1, keepdims=True)
    assert np.allclose(vn.cpu().numpy(), acc, atol=2e-4)
    assert (vn.cpu().numpy() * (vh / np.linalg.norm(vh, axis=1, keepdims=True))).sum(1).min() > 0.99      # outward on a sphere
    # colours
    params = init_params(dict(NGP_DEFAULTS), seed=5, grid_scale=0.5)
    pipe = NGPPipeline(max_rays=256, capacity=1 << 16, params={k: t.clone() for k, t in params.items()})
    comp = pipe.mesh_components(v, f)
    assert comp['vert_colors'].shape == (v.shape[0], 3) and comp['face_colors'].shape == (f.shape[0], 3)
    sig, rgb = pipe.forward_pts_dir(v, -vn)
    # vertex normals are summed with float atomics: two evaluations differ in the last bits, and so do the view directions
    assert torch.allclose(rgb, comp['vert_colors'], atol=2e-3)
    assert torch.allclose(sig, pipe.forward_pts(v), rtol=2e-3, atol=1e-4)
    orc = NGPOracle({k: t.numpy() for k, t in params.items()}, NGP_CONFIG)
    unit = ((v.cpu().numpy() - orc.aabb[0]) / (orc.aabb[1] - orc.aabb[0])).astype(np.float32)
    s_ref, c_ref, _ = orc.field(unit, (-vn).cpu().numpy())
    assert np.allclose(rgb.cpu().numpy(), c_ref.numpy(), atol=4e-3)
    assert np.allclose(sig.cpu().numpy(), s_ref.numpy(), rtol=2e-2, atol=1e-3)


def test_slab_sharded_extraction_tiles_the_whole_mesh():
    """mesh extraction sharded by x slabs (bench config 'mesh', dist.shard_range): the cells of the slabs tile the lattice, so the
    slab meshes together have the faces of the whole-lattice mesh, and every slab vertex is a vertex of the whole mesh"""
    from arcnerf_b200 import dist as adist
    from arcnerf_b200.ngp import NGP_DEFAULTS, NGPPipeline, init_params
    pipe = NGPPipeline(max_rays=256, capacity=1 << 16, params=init_params(dict(NGP_DEFAULTS), seed=7, grid_scale=1.0))
    n = 40
    level = float(pipe.density_grid(n).median())
    v, f, _ = pipe.extract_mesh(n, level=level)
    from scipy.spatial import cKDTree
    whole = cKDTree(v.cpu().numpy())
    n_faces, world = 0, 3
    for rank in range(world):
        x0, x1 = adist.shard_range(n, rank, world)
        vs_, fs_, _ = pipe.extract_mesh(n, level=level, x_range=(x0, x1))
        n_faces += fs_.shape[0]
        if vs_.shape[0]:                     # the slab's x positions are offset after the kernel: equal to fp32 rounding
            assert whole.query(vs_.cpu().numpy())[0].max() < 1e-5
        if fs_.shape[0]:
            assert int(fs_.max()) < vs_.shape[0]
    assert n_faces == f.shape[0] and f.shape[0] > 0


def _look_at(eye, target=(0.0, 0.0, 0.0)):
    eye, target = np.asarray(eye, np.float64), np.asarray(target, np.float64)
    fwd = target - eye
    fwd /= np.linalg.norm(fwd)
    right = np.cross(fwd, np.array([0.0, 0.0, 1.0]))
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    c2w = np.eye(4)
    c2w[:3, 0], c2w[:3, 1], c2w[:3, 2], c2w[:3, 3] = right, down, fwd, eye
    return np.linalg.inv(c2w).astype(np.float32)


def test_mesh_rasteriser_equals_restatement():
    """csrc/raster.cu against oracle/raster.py on an extracted sphere mesh seen from three cameras (one inside the mesh, so that
    faces behind the near plane are skipped): face-id buffer, depth and colours"""
    from arcnerf_b200.plugin.mesh import extract_mesh, render_mesh_images
    from oracle import raster as orr
    vol, vs, vl = sphere_volume(24, 0.6)
    v, f, n = extract_mesh(vol, 50.0, vs, vl)
    vc = (0.5 + 0.5 * n).astype(np.float32)
    H, W = 48, 64
    K = np.array([[70.0, 0, W / 2], [0, 70.0, H / 2], [0, 0, 1]], np.float32)
    w2c = np.stack([_look_at((2.0, 0.3, 0.4)), _look_at((-0.5, 1.8, -0.7)), _look_at((0.1, 0.0, 0.05), (1.0, 0.2, 0.0))])
    img, fid, depth = render_mesh_images(v, f, vc, None, n, None, H, W, w2c, K, return_buffers=True)
    r_img, r_fid, r_depth = orr.raster_mesh(v, f, w2c, K, H, W, vert_colors=vc)
    assert img.shape == (3, H, W, 3) and img.dtype == np.uint8
    assert np.array_equal(fid.cpu().numpy(), r_fid) and (r_fid[0] >= 0).sum() > 200 and (r_fid[0] < 0).sum() > 200
    assert np.array_equal(depth.cpu().numpy().view(np.uint32), r_depth.view(np.uint32))
    assert np.abs(img.astype(np.int32) - r_img.astype(np.int32)).max() <= 1           # fp32 colour blend -> the uint8 rounding edge
    # face colours and silhouette
    fc = np.random.RandomState(0).uniform(0, 1, (len(f), 3)).astype(np.float32)
    img2 = render_mesh_images(v, f, None, fc, None, None, H, W, w2c[:1], K)
    assert np.array_equal(img2, orr.raster_mesh(v, f, w2c[:1], K, H, W, face_colors=fc)[0])
    sil = render_mesh_images(v, f, vc, None, None, None, H, W, w2c[:1], K, sil_mode=True)
    assert np.array_equal(sil[0, :, :, 0] > 0, r_fid[0] >= 0)
    # empty mesh: background only
    empty = render_mesh_images(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.int32), None, None, None, None, 8, 8, w2c[:1], K)
    assert (empty == 255).all()
