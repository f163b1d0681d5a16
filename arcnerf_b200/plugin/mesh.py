# -*- coding: utf-8 -*-
"""Mesh extraction on the device: drop-ins for arcnerf/geometry/mesh.py:12-47 extract_mesh (same arguments
and return order; skimage's host marching cubes replaced by csrc/mcubes.cu) and for the lattice of voxel centres that
run_infer_volume feeds the model (arcnerf/eval/infer_func.py:448-500, Volume.get_volume_pts)."""
import ctypes as C

import numpy as np
import torch

from .. import _lib


def volume_pts(lo, hi, n_grid, x0=0, x1=None, device='cuda', out=None):
    """voxel centres of x slabs [x0, x1) of the n_grid^3 lattice over [lo, hi] -> ((x1 - x0) * n_grid^2, 3) fp32 on the device"""
    x1 = n_grid if x1 is None else x1
    n = (x1 - x0) * n_grid * n_grid
    if out is None:
        out = torch.empty((n, 3), dtype=torch.float32, device=device)
    assert out.is_cuda and out.is_contiguous() and out.numel() >= 3 * n
    lo_c = (C.c_float * 3)(*[float(v) for v in lo])
    hi_c = (C.c_float * 3)(*[float(v) for v in hi])
    with torch.cuda.device(out.device):
        _lib.call('arcb200_volume_pts', lo_c, hi_c, int(n_grid), int(x0), int(x1), out.data_ptr(), _lib.current_stream())
    return out[:n] if out.dim() == 2 else out


def extract_mesh(sigma, level, voxel_size, volume_len, grad_dir='descent', with_normals=True):
    """sigma (nx, ny, nz): CUDA tensor (results stay on the device) or numpy array (results come back as numpy, like the
    reference).  Returns verts (V, 3) fp32, faces (F, 3) int32 with outward winding, vert_normals (V, 3) unit, outward."""
    assert grad_dir in ['descent', 'ascent'], 'Invalid grad_dir, only descent/ascent'
    as_numpy = isinstance(sigma, np.ndarray)
    vol = torch.from_numpy(np.ascontiguousarray(sigma, np.float32)).cuda() if as_numpy else sigma.contiguous().float()
    if not vol.is_cuda or vol.dim() != 3:
        raise RuntimeError('sigma must be a (nx, ny, nz) volume on a CUDA device (or a numpy array)')
    nx, ny, nz = vol.shape
    dev, ascent = vol.device, int(grad_dir == 'ascent')
    sizes = (C.c_int64 * 3)()
    _lib.call('arcb200_mc_scratch_sizes', nx, ny, nz, sizes)
    with torch.cuda.device(dev):
        st = _lib.current_stream()
        pblk = torch.empty(sizes[0], dtype=torch.int32, device=dev)
        cblk = torch.empty(sizes[1], dtype=torch.int32, device=dev)
        _lib.call('arcb200_mc_count', vol.data_ptr(), nx, ny, nz, float(level), ascent, pblk.data_ptr(), cblk.data_ptr(), st)
        n_v, n_f = int(pblk[-1]), int(cblk[-1])                              # the one host sync: sizes of the outputs
        verts = torch.empty((n_v, 3), dtype=torch.float32, device=dev)
        normals = torch.empty((n_v, 3), dtype=torch.float32, device=dev) if with_normals else None
        faces = torch.empty((n_f, 3), dtype=torch.int32, device=dev)
        vbase = torch.empty(sizes[2], dtype=torch.int32, device=dev)
        spacing = (C.c_float * 3)(*[float(v) for v in voxel_size])
        offset = (C.c_float * 3)(*[(float(volume_len[a]) - float(voxel_size[a])) / 2.0 for a in range(3)])
        _lib.call('arcb200_mc_emit', vol.data_ptr(), nx, ny, nz, float(level), ascent, spacing, offset, pblk.data_ptr(),
                  cblk.data_ptr(), vbase.data_ptr(), verts.data_ptr(), normals.data_ptr() if with_normals else None, n_v,
                  faces.data_ptr(), n_f, st)
    if as_numpy:
        return verts.cpu().numpy(), faces.cpu().numpy(), normals.cpu().numpy() if with_normals else None
    return verts, faces, normals


def get_mesh_components(verts, faces):
    """get_normals + get_face_centers (arcnerf/geometry/mesh.py:88-125) on the device: verts (V, 3) fp32, faces (F, 3) int32
    CUDA tensors -> vert_normals (V, 3), face_normals (F, 3), face_centers (F, 3)"""
    if not (verts.is_cuda and faces.is_cuda):
        raise RuntimeError('verts and faces must be CUDA tensors')
    verts, faces = verts.contiguous().float(), faces.contiguous().to(torch.int32)
    n_v, n_f = verts.shape[0], faces.shape[0]
    vn = torch.empty((n_v, 3), dtype=torch.float32, device=verts.device)
    fn = torch.empty((n_f, 3), dtype=torch.float32, device=verts.device)
    fc = torch.empty((n_f, 3), dtype=torch.float32, device=verts.device)
    with torch.cuda.device(verts.device):
        _lib.call('arcb200_mesh_components', verts.data_ptr(), faces.data_ptr(), n_v, n_f, fn.data_ptr(), fc.data_ptr(), vn.data_ptr(),
                  _lib.current_stream())
    return vn, fn, fc


def render_mesh_images(verts, faces, vert_colors, face_colors, vert_normals, face_normals, H, W, w2c, intrinsic, backend='b200',
                       single_image_mode=False, device='cuda', sil_mode=False, sil_hard=True, return_buffers=False):
    """arcnerf/geometry/mesh.py:174-240 render_mesh_images with the same arguments; `backend` is ignored (the one rasteriser of
    this build, csrc/raster.cu, replaces both open3d and pytorch3d; normals are accepted and unused: the colours are unlit).
    Returns images (N, H, W, 3) uint8 as numpy (sil_mode: the silhouette, covered pixels white on black like the reference's
    hard silhouette); return_buffers=True adds the face-id and depth buffers (device tensors)."""
    dev = torch.device(device if str(device) != 'gpu' else 'cuda')
    t = lambda a, dt: (a.to(dev, dt) if torch.is_tensor(a) else torch.as_tensor(np.asarray(a), dtype=dt, device=dev)).contiguous()
    v, f = t(verts, torch.float32), t(faces, torch.int32)
    cams = t(w2c, torch.float32).reshape(-1, 4, 4).contiguous()
    n_views = cams.shape[0]
    vc = t(vert_colors, torch.float32) if (vert_colors is not None and not sil_mode) else None
    fc = t(face_colors, torch.float32) if (face_colors is not None and vert_colors is None and not sil_mode) else None
    k = (C.c_float * 9)(*[float(x) for x in np.asarray(intrinsic.cpu() if torch.is_tensor(intrinsic) else intrinsic, np.float64).reshape(-1)])
    nbytes = C.c_int64()
    _lib.call('arcb200_raster_scratch_bytes', n_views, int(H), int(W), C.byref(nbytes))
    scratch = torch.empty(max(1, nbytes.value), dtype=torch.uint8, device=dev)
    image = torch.empty((n_views, H, W, 3), dtype=torch.uint8, device=dev)
    fid = torch.empty((n_views, H, W), dtype=torch.int32, device=dev)
    depth = torch.empty((n_views, H, W), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _lib.call('arcb200_raster_mesh', v.data_ptr(), f.data_ptr(), _lib.ptr(vc), _lib.ptr(fc), cams.data_ptr(), k, n_views, int(H), int(W),
                  1e-6, 0.0 if sil_mode else 1.0, scratch.data_ptr(), image.data_ptr(), fid.data_ptr(), depth.data_ptr(), v.shape[0],
                  f.shape[0], _lib.current_stream())
    if sil_mode:
        image = ((fid >= 0).to(torch.uint8) * 255)[..., None].expand(-1, -1, -1, 3).contiguous()
    out = image.cpu().numpy()
    return (out, fid, depth) if return_buffers else out
