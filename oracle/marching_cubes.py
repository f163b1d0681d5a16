"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): numpy restatement of the mesh-extraction path, SURVEY.md §8 f1.

  * volume_pts: Volume.cal_volume_pts, arcnerf/geometry/volume.py:225-232 (torch.linspace per axis, stated here with its
    symmetric, fused-multiply-add evaluation rule; bit-exact against tests/golden/volume_golden.npz);
  * extract_mesh: arcnerf/geometry/mesh.py:12-47.  The reference calls skimage.measure.marching_cubes (Lewiner), a third-party
    dependency that is neither vendored in /root/reference nor installed in this image (scikit-image, unpinned in the reference's
    requirements).  PARITY UNPINNED for vertex/face order and for the choice made in ambiguous cells: this restatement is the
    classic 256-case marching cubes with the table of arcnerf_b200/tools/gen_mc_table.py (derived by contour tracing, ambiguous
    faces cut off inside corners), linear interpolation of the level crossing on lattice edges (what skimage does as well),
    central-difference vertex normals.  What the tests pin instead: this restatement against closed-form surfaces (sphere area,
    volume, vertex distance), closedness and orientation of the mesh; the CUDA kernels against this restatement bit for bit.
"""
import numpy as np

from arcnerf_b200.tools import gen_mc_table as T

F32 = np.float32


def volume_pts(lo, hi, n_grid, x0=0, x1=None):
    x1 = n_grid if x1 is None else x1
    lo, hi = np.asarray(lo, F32), np.asarray(hi, F32)
    axes = []
    for a in range(3):
        v = F32((hi[a] - lo[a]) / F32(n_grid))
        s, e = F32(float(lo[a]) + 0.5 * float(v)), F32(float(hi[a]) - 0.5 * float(v))
        step = F32((e - s) / F32(n_grid - 1)) if n_grid > 1 else F32(0)
        i = np.arange(n_grid)
        # one fused multiply-add per value (ATen's CPU linspace): exact product in float64, one rounding
        axes.append(np.where(i < n_grid // 2, np.float64(s) + np.float64(step) * i,
                             np.float64(e) - np.float64(step) * (n_grid - 1 - i)).astype(F32))
    g = np.stack(np.meshgrid(axes[0][x0:x1], axes[1], axes[2], indexing='ij'), -1)
    return g.reshape(-1, 3)


def _lattice_grad(vol, axis):
    """central differences (one-sided on the border), index units"""
    n = vol.shape[axis]
    idx = np.arange(n)
    lo, hi = np.maximum(idx - 1, 0), np.minimum(idx + 1, n - 1)
    d = (np.take(vol, hi, axis) - np.take(vol, lo, axis)).astype(F32)
    shape = [1, 1, 1]
    shape[axis] = n
    return (d / (hi - lo).astype(F32).reshape(shape)).astype(F32)


def extract_mesh(sigma, level, voxel_size, volume_len, grad_dir='descent', with_normals=True):
    """-> verts (V, 3) f32, faces (F, 3) i32 (normals by the right-hand rule point outside), vert_normals (V, 3) f32"""
    assert grad_dir in ['descent', 'ascent'], 'Invalid grad_dir, only descent/ascent'
    vol = np.ascontiguousarray(sigma, F32)
    nx, ny, nz = vol.shape
    level = F32(level)
    inside = (vol < level) if grad_dir == 'ascent' else (vol > level)
    spacing = np.asarray(voxel_size, F32)
    offset = np.asarray([(float(volume_len[a]) - float(voxel_size[a])) / 2.0 for a in range(3)], F32)

    # --- vertices: one per crossing +axis edge, in lattice order (point major, axis minor)
    cross = np.zeros((nx, ny, nz, 3), bool)
    cross[:-1, :, :, 0] = inside[1:] != inside[:-1]
    cross[:, :-1, :, 1] = inside[:, 1:] != inside[:, :-1]
    cross[:, :, :-1, 2] = inside[:, :, 1:] != inside[:, :, :-1]
    n_at = cross.sum(-1).reshape(-1)
    vbase = np.concatenate([[0], np.cumsum(n_at)])[:-1].astype(np.int64)
    pi, pj, pk, ax = np.nonzero(cross)                                       # already in (point, axis) order
    q = [pi + (ax == 0), pj + (ax == 1), pk + (ax == 2)]
    f0, f1 = vol[pi, pj, pk], vol[q[0], q[1], q[2]]
    with np.errstate(all='ignore'):
        t = ((level - f0) / (f1 - f0)).astype(F32)
    pos = np.stack([pi, pj, pk], -1).astype(F32)
    pos[np.arange(len(ax)), ax] += t
    verts = (pos * spacing - offset).astype(F32)
    normals = None
    if with_normals:
        g = [_lattice_grad(vol, a) for a in range(3)]
        n = np.stack([g[b][pi, pj, pk] + t * (g[b][q[0], q[1], q[2]] - g[b][pi, pj, pk]) for b in range(3)], -1).astype(F32)
        n = (n / spacing).astype(F32) * F32(1.0 if grad_dir == 'ascent' else -1.0)
        ln = np.sqrt((n[:, 0] * n[:, 0] + n[:, 1] * n[:, 1]).astype(F32) + n[:, 2] * n[:, 2]).astype(F32) + F32(1e-8)
        normals = (n / ln[:, None]).astype(F32)

    # --- faces: cells in linear order (z fastest), the table's triangles in order
    cs = np.zeros((nx - 1, ny - 1, nz - 1), np.int32)
    for c, (dx, dy, dz) in enumerate(T.CORNERS):
        cs |= inside[dx:nx - 1 + dx, dy:ny - 1 + dy, dz:nz - 1 + dz].astype(np.int32) << c
    rank = np.cumsum(cross, -1) - cross                                      # axis rank among a point's crossing edges
    vid_point = vbase.reshape(nx, ny, nz)
    ci, cj, ck = np.nonzero((cs != 0) & (cs != 255))
    faces = []
    cases = cs[ci, cj, ck]
    order_cell = np.arange(len(cases))
    out_cell, out_tri, out_f = [], [], []
    for case in np.unique(cases):
        sel = np.nonzero(cases == case)[0]
        for ti, tri in enumerate(T.TABLE[case]):
            cols = []
            for e in tri:
                _, _, axis, (ox, oy, oz) = T.EDGES[e]
                a, b, c = ci[sel] + ox, cj[sel] + oy, ck[sel] + oz
                cols.append(vid_point[a, b, c] + rank[a, b, c, axis])
            out_f.append(np.stack(cols, -1))
            out_cell.append(order_cell[sel])
            out_tri.append(np.full(len(sel), ti))
    if out_f:
        f = np.concatenate(out_f)
        o = np.lexsort((np.concatenate(out_tri), np.concatenate(out_cell)))
        faces = f[o].astype(np.int32)
    else:
        faces = np.zeros((0, 3), np.int32)
    return verts, faces, normals
