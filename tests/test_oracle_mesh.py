"""CPU checks of the mesh-extraction oracle (oracle/marching_cubes.py) and of the generated case table: the reference's Volume
lattice (golden), closed-form surfaces, and structural properties of all 256 cases."""
import os

import numpy as np
import pytest

from arcnerf_b200.tools import gen_mc_table as T
from oracle import marching_cubes as M
from mesh_cases import blobs_volume, closed_and_oriented, sphere_volume

GOLD = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'volume_golden.npz'))


@pytest.mark.parametrize('tag', ['cube16', 'box9', 'one', 'box67'])
def test_volume_pts_bit_exact_with_reference_volume(tag):
    lo, hi, n = GOLD['%s_lo' % tag], GOLD['%s_hi' % tag], int(GOLD['%s_n' % tag])
    got, ref = M.volume_pts(lo, hi, n), GOLD['%s_pts' % tag]
    assert got.shape == ref.shape and np.array_equal(got.view(np.uint32), ref.view(np.uint32))


def test_case_table_structure():
    assert T.MAX_TRIS == 5 and len(T.TABLE) == 256 and not T.TABLE[0] and not T.TABLE[255]
    for case, tris in enumerate(T.TABLE):
        crossing = {e for e, (a, b, _, _) in enumerate(T.EDGES) if ((case >> a) & 1) != ((case >> b) & 1)}
        used = {e for t in tris for e in t}
        assert used == crossing, case                                       # every crossing edge is a vertex, no others
        assert all(len(set(t)) == 3 for t in tris)
    # the committed header is the generator's output
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        T.write_header(os.path.join(d, 'h'))
        here = os.path.join(os.path.dirname(T.__file__), '..', 'csrc', 'mc_table.cuh')
        assert open(os.path.join(d, 'h')).read() == open(here).read()


def test_sphere_matches_closed_form():
    n, r = 64, 0.6
    vol, vs, vl = sphere_volume(n, r)
    v, f, nr = M.extract_mesh(vol, 50.0, vs, vl)
    closed_and_oriented(v, f)
    d = np.linalg.norm(v, axis=1)
    assert abs(d - r).max() < 3e-3
    a, b, c = v[f[:, 0]], v[f[:, 1]], v[f[:, 2]]
    cr = np.cross(b - a, c - a)
    assert abs(0.5 * np.linalg.norm(cr, axis=1).sum() / (4 * np.pi * r * r) - 1) < 4e-3
    assert abs((a * cr).sum() / 6 / (4 / 3 * np.pi * r ** 3) - 1) < 4e-3    # positive volume: outward winding
    assert (nr * (v / d[:, None])).sum(1).min() > 0.999
    assert len(v) - 3 * len(f) // 2 + len(f) == 2


def test_noise_volume_is_closed_for_every_case_and_direction():
    rng = np.random.default_rng(5)
    vol = np.zeros((20, 20, 20), np.float32)
    vol[2:-2, 2:-2, 2:-2] = rng.random((16, 16, 16), dtype=np.float32)
    v, f, _ = M.extract_mesh(vol, 0.5, (1., 1., 1.), (20., 20., 20.))
    closed_and_oriented(v, f, manifold=False)
    v2, f2, _ = M.extract_mesh(-vol, -0.5, (1., 1., 1.), (20., 20., 20.), 'ascent')
    assert np.array_equal(v, v2) and np.array_equal(f, f2)                  # same inside set -> same mesh


def test_vertices_lie_on_the_level_set_of_the_trilinear_edges():
    vol, vs, vl, level = blobs_volume((24, 18, 30), seed=1)
    v, f, _ = M.extract_mesh(vol, level, vs, vl)
    # back to index space; exactly one coordinate is fractional and the linear interpolant there equals the level
    idx = (v + (np.asarray(vl) - np.asarray(vs)) / 2) / np.asarray(vs)
    frac = np.abs(idx - np.round(idx))
    ax = frac.argmax(1)
    lo = np.floor(idx + 1e-4).astype(int)
    lo[np.arange(len(v)), ax] = np.floor(idx[np.arange(len(v)), ax]).astype(int)
    lo = np.minimum(lo, np.asarray(vol.shape) - 1)
    hi = lo.copy()
    hi[np.arange(len(v)), ax] = np.minimum(hi[np.arange(len(v)), ax] + 1, np.asarray(vol.shape)[ax] - 1)
    t = idx[np.arange(len(v)), ax] - lo[np.arange(len(v)), ax]
    val = vol[lo[:, 0], lo[:, 1], lo[:, 2]] * (1 - t) + vol[hi[:, 0], hi[:, 1], hi[:, 2]] * t
    assert np.abs(val - level).max() < 2e-2 * np.abs(vol).max()


def test_raster_restatement_covers_the_projected_sphere():
    """oracle/raster.py on an extracted sphere: the covered pixels form the projected disc, depth runs from the near pole to the
    silhouette, colours are the interpolated vertex colours, a camera inside the mesh skips the faces behind its near plane"""
    from oracle import raster as orr
    n, r = 24, 0.6
    vol, vs, vl = sphere_volume(n, r)
    v, f, nr = M.extract_mesh(vol, 50.0, vs, vl)
    H, W, foc, dist = 48, 64, 70.0, 2.0
    K = np.array([[foc, 0, W / 2], [0, foc, H / 2], [0, 0, 1]], np.float32)
    w2c = np.eye(4, dtype=np.float32)
    w2c[2, 3] = dist                                          # camera on the -z axis looking along +z
    img, fid, depth = orr.raster_mesh(v, f, w2c[None], K, H, W, vert_colors=(0.5 + 0.5 * nr).astype(np.float32))
    cov = fid[0] >= 0
    # silhouette radius of a sphere seen from distance d: f r / sqrt(d^2 - r^2)
    rad = foc * r / np.sqrt(dist * dist - r * r)
    assert abs(cov.sum() / (np.pi * rad * rad) - 1) < 0.05
    assert abs(depth[0][cov].min() - (dist - r)) < 0.02 and depth[0][cov].max() < dist
    assert (img[0][~cov] == 255).all()
    centre = img[0, H // 2, W // 2].astype(np.float32) / 255.0      # normal at the near pole is -z: colour (0.5, 0.5, 0)
    assert np.abs(centre - np.array([0.5, 0.5, 0.0])).max() < 0.08
    inside_cam = np.eye(4, dtype=np.float32)
    _, fid2, depth2 = orr.raster_mesh(v, f, inside_cam[None], K, H, W)
    assert (fid2[0] >= 0).all() and abs(depth2[0, H // 2, W // 2] - r) < 0.02      # only the far half is in front of the camera
