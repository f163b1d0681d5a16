"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): numpy restatement of the mesh preview rasteriser (csrc/raster.cu), SURVEY.md
§8 f4.  The reference renders mesh previews through pytorch3d (arcnerf/render/render_pytorch3d.py, arcnerf/geometry/mesh.py:
174-240), which is not in this image: PARITY UNPINNED against the reference.  What this pins is the kernel: the same pinhole
projection, pixel-centre coverage test, perspective-correct depth and colour, nearest face with ties to the lower face id, one
rounded fp32 operation at a time."""
import numpy as np

F32 = np.float32


def _edge(ax, ay, bx, by, px, py):
    return ((bx - ax) * (py - ay)).astype(F32) - ((by - ay) * (px - ax)).astype(F32)


def raster_mesh(verts, faces, w2c, intrinsic, H, W, vert_colors=None, face_colors=None, near=1e-6, bkg=1.0):
    """-> image (N, H, W, 3) uint8, face_id (N, H, W) int32, depth (N, H, W) f32"""
    verts, w2c, K = np.asarray(verts, F32), np.asarray(w2c, F32).reshape(-1, 4, 4), np.asarray(intrinsic, F32)
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    n_views = w2c.shape[0]
    image = np.empty((n_views, H, W, 3), np.uint8)
    fid = -np.ones((n_views, H, W), np.int32)
    zbuf = np.full((n_views, H, W), np.inf, F32)
    col = np.full((n_views, H, W, 3), F32(bkg), F32)
    near = F32(near)
    for v in range(n_views):
        R = w2c[v]
        x = (((R[0, 0] * verts[:, 0]).astype(F32) + (R[0, 1] * verts[:, 1]).astype(F32)).astype(F32) + (R[0, 2] * verts[:, 2]).astype(F32)).astype(F32) + R[0, 3]
        y = (((R[1, 0] * verts[:, 0]).astype(F32) + (R[1, 1] * verts[:, 1]).astype(F32)).astype(F32) + (R[1, 2] * verts[:, 2]).astype(F32)).astype(F32) + R[1, 3]
        z = (((R[2, 0] * verts[:, 0]).astype(F32) + (R[2, 1] * verts[:, 1]).astype(F32)).astype(F32) + (R[2, 2] * verts[:, 2]).astype(F32)).astype(F32) + R[2, 3]
        with np.errstate(all='ignore'):
            u = ((fx * x).astype(F32) / z).astype(F32) + cx
            w = ((fy * y).astype(F32) / z).astype(F32) + cy
        for f, (i0, i1, i2) in enumerate(np.asarray(faces)):
            zz = np.array([z[i0], z[i1], z[i2]], F32)
            if not (zz > near).all():
                continue
            uu, vv = np.array([u[i0], u[i1], u[i2]], F32), np.array([w[i0], w[i1], w[i2]], F32)
            area = _edge(uu[0], vv[0], uu[1], vv[1], uu[2], vv[2])
            if area == 0 or not np.isfinite(area):
                continue
            if not (uu.max() >= 0 and vv.max() >= 0 and uu.min() <= W and vv.min() <= H):
                continue
            c0, c1 = max(0, int(np.floor(uu.min() - F32(0.5)))), min(W - 1, int(np.ceil(uu.max() - F32(0.5))))
            r0, r1 = max(0, int(np.floor(vv.min() - F32(0.5)))), min(H - 1, int(np.ceil(vv.max() - F32(0.5))))
            if c1 < c0 or r1 < r0:
                continue
            px, py = np.meshgrid(np.arange(c0, c1 + 1, dtype=F32) + F32(0.5), np.arange(r0, r1 + 1, dtype=F32) + F32(0.5))
            w0 = _edge(uu[1], vv[1], uu[2], vv[2], px, py)
            w1 = _edge(uu[2], vv[2], uu[0], vv[0], px, py)
            w2 = _edge(uu[0], vv[0], uu[1], vv[1], px, py)
            inside = ((w0 >= 0) & (w1 >= 0) & (w2 >= 0)) if area > 0 else ((w0 <= 0) & (w1 <= 0) & (w2 <= 0))
            b0, b1, b2 = (w0 / area).astype(F32), (w1 / area).astype(F32), (w2 / area).astype(F32)
            q0, q1, q2 = (b0 / zz[0]).astype(F32), (b1 / zz[1]).astype(F32), (b2 / zz[2]).astype(F32)
            s = ((q0 + q1).astype(F32) + q2).astype(F32)
            with np.errstate(all='ignore'):
                depth = (F32(1.0) / s).astype(F32)
            # nearest wins; equal depth: the lower face id, which is the one already stored (faces are visited in order)
            take = inside & (depth > near) & (depth < zbuf[v, r0:r1 + 1, c0:c1 + 1])
            if not take.any():
                continue
            zb = zbuf[v, r0:r1 + 1, c0:c1 + 1]
            zb[take] = depth[take]
            fid[v, r0:r1 + 1, c0:c1 + 1][take] = f
            if face_colors is not None:
                c = np.broadcast_to(np.asarray(face_colors, F32)[f], take.shape + (3, ))
            elif vert_colors is not None:
                vc = np.asarray(vert_colors, F32)
                c = ((((q0[..., None] * vc[i0]).astype(F32) + (q1[..., None] * vc[i1]).astype(F32)).astype(F32) +
                      (q2[..., None] * vc[i2]).astype(F32)).astype(F32) / s[..., None]).astype(F32)
            else:
                c = np.zeros(take.shape + (3, ), F32)
            col[v, r0:r1 + 1, c0:c1 + 1][take] = c[take]
        image[v] = (np.clip(col[v], 0, 1) * F32(255.0) + F32(0.5)).astype(np.uint8)
    depth_out = np.where(np.isfinite(zbuf), zbuf, F32(0)).astype(F32)
    return image, fid, depth_out
