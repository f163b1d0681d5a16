"""Seeded synthetic inputs shaped like the reference's data: Blender-shaped pinhole
rays, occupancy volumes in both of the reference's storage forms, and tcnn-layout random parameters.

There are no datasets on the box; everything here is numpy + a fixed seed so that the CUDA path, the
CPU oracle and the committed golden fixtures all see identical bytes.
"""
import math

import numpy as np

BLENDER_ANGLE_X = 0.6911112070083618  # NeRF-synthetic `camera_angle_x` (lego)


def look_at_pose(cam_loc, target=(0.0, 0.0, 0.0), up=(0.0, 0.0, 1.0)):
    """c2w (4,4) of a pinhole camera at cam_loc looking at target (x right, y down, z forward)."""
    cam_loc = np.asarray(cam_loc, np.float64)
    fwd = np.asarray(target, np.float64) - cam_loc
    fwd /= np.linalg.norm(fwd)
    right = np.cross(fwd, np.asarray(up, np.float64))
    if np.linalg.norm(right) < 1e-8:
        right = np.cross(fwd, np.array([0.0, 1.0, 0.0]))
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    c2w = np.eye(4)
    c2w[:3, 0], c2w[:3, 1], c2w[:3, 2], c2w[:3, 3] = right, down, fwd, cam_loc
    return c2w


def sphere_poses(n, radius=3.0, seed=0, v_range=(0.15, 0.85)):
    """n camera poses on the upper part of a sphere of `radius`, all looking at the origin."""
    rng = np.random.RandomState(seed)
    u = rng.uniform(0.0, 2.0 * math.pi, n)
    v = rng.uniform(v_range[0], v_range[1], n) * (math.pi / 2)
    locs = np.stack([radius * np.cos(u) * np.cos(v), radius * np.sin(u) * np.cos(v), radius * np.sin(v)], -1)
    return np.stack([look_at_pose(p) for p in locs], 0)


def camera_rays(W, H, c2w, angle_x=BLENDER_ANGLE_X, center_pixel=True, index=None):
    """Rays of one pinhole view, flattened column-major (pixel (i, j) -> i*H + j) like the reference's
    ray_helper.get_rays(wh_order=True) (arcnerf/render/ray_helper.py:12-119); directions normalised.
    focal = 0.5*W/tan(0.5*angle_x) (arcnerf/datasets/nerf_dataset.py:165-174)."""
    focal = 0.5 * W / math.tan(0.5 * angle_x)
    if index is None:
        i, j = np.meshgrid(np.arange(W, dtype=np.float64), np.arange(H, dtype=np.float64), indexing='ij')
        i, j = i.reshape(-1), j.reshape(-1)
    else:
        index = np.asarray(index)
        i, j = (index // H).astype(np.float64), (index % H).astype(np.float64)
    if center_pixel:
        i, j = i + 0.5, j + 0.5
    cam = np.stack([(i - W / 2.0) / focal, (j - H / 2.0) / focal, np.ones_like(i)], -1)
    d = cam @ c2w[:3, :3].T
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    o = np.broadcast_to(c2w[:3, 3], d.shape)
    return np.ascontiguousarray(o, np.float32), np.ascontiguousarray(d, np.float32)


def train_ray_batch(n_rays, seed=0, W=800, H=800, n_poses=100, radius=3.0):
    """A random batch drawn from a pool of n_poses Blender-shaped views (the reference concatenates the
    rays of all training images and draws by randperm: arcnerf_trainer.py:188-219, pipeline.py:204-220)."""
    rng = np.random.RandomState(seed)
    poses = sphere_poses(n_poses, radius, seed=1234)
    which = rng.randint(0, n_poses, n_rays)
    pix = rng.randint(0, W * H, n_rays)
    o = np.empty((n_rays, 3), np.float32)
    d = np.empty((n_rays, 3), np.float32)
    for p in np.unique(which):
        sel = np.nonzero(which == p)[0]
        o[sel], d[sel] = camera_rays(W, H, poses[p], index=pix[sel])
    return o, d


def aabb_of_side(side=2.0, origin=(0.0, 0.0, 0.0)):
    """(2,3) [min xyz; max xyz] of an axis-aligned cube."""
    c = np.asarray(origin, np.float32)
    return np.stack([c - side / 2.0, c + side / 2.0], 0).astype(np.float32)


def _voxel_centres(n, side=2.0):
    ax = (np.arange(n, dtype=np.float32) + 0.5) / n * side - side / 2.0
    return np.meshgrid(ax, ax, ax, indexing='ij')


def occupancy_volume(kind='lego', n=128, seed=0, side=2.0):
    """(n,n,n) bool occupancy, indexed [x][y][z] like Volume's bitfield (flat index x*n*n + y*n + z,
    arcnerf/ops/include/volume_func.h:59-68).

    kind: 'sphere'  - analytic ball of radius 0.5*side*0.5 (the FakeGeoNet density, tests_fg_model.py:293-309)
          'lego'    - union of boxes filling ~10-15 % of the volume (a trained-NGP-like pattern)
          'core50'  - cubic core at 50 % random occupancy (tests_models/__init__.py:216-241)
          'full' / 'empty'
    """
    X, Y, Z = _voxel_centres(n, side)
    if kind == 'full':
        return np.ones((n, n, n), bool)
    if kind == 'empty':
        return np.zeros((n, n, n), bool)
    if kind == 'sphere':
        return (X * X + Y * Y + Z * Z) < (0.25 * side) ** 2
    if kind == 'core50':
        rng = np.random.RandomState(seed)
        core = (np.abs(X) < 0.25 * side) & (np.abs(Y) < 0.25 * side) & (np.abs(Z) < 0.25 * side)
        return core & (rng.uniform(size=(n, n, n)) < 0.5)
    if kind == 'lego':
        rng = np.random.RandomState(seed)
        occ = np.zeros((n, n, n), bool)
        h = 0.5 * side
        # a base plate, a body, and a scatter of small bricks
        boxes = [((-0.7, -0.4, -0.55), (0.7, 0.4, -0.40)), ((-0.45, -0.3, -0.40), (0.35, 0.3, 0.05)),
                 ((0.1, -0.15, 0.05), (0.6, 0.15, 0.45))]
        for _ in range(24):
            c = rng.uniform(-0.6, 0.6, 3)
            s = rng.uniform(0.04, 0.16, 3)
            boxes.append((tuple(c - s), tuple(c + s)))
        for lo, hi in boxes:
            occ |= ((X > lo[0] * h) & (X < hi[0] * h) & (Y > lo[1] * h) & (Y < hi[1] * h) &
                    (Z > lo[2] * h) & (Z < hi[2] * h))
        return occ
    raise ValueError(kind)


def _spread3(v):
    v = v.astype(np.uint32)
    v = (v * np.uint32(0x00010001)) & np.uint32(0xFF0000FF)
    v = (v * np.uint32(0x00000101)) & np.uint32(0x0F00F00F)
    v = (v * np.uint32(0x00000011)) & np.uint32(0xC30C30C3)
    v = (v * np.uint32(0x00000005)) & np.uint32(0x49249249)
    return v


def morton_index(n):
    """(n,n,n) uint32 Morton code of voxel [x][y][z] (volume_func.h:142-157)."""
    ax = np.arange(n, dtype=np.uint32)
    X, Y, Z = np.meshgrid(ax, ax, ax, indexing='ij')
    return _spread3(X) | (_spread3(Y) << np.uint32(1)) | (_spread3(Z) << np.uint32(2))


def volume_to_bitfield(vol):
    """(n,n,n) bool -> (n**3/8,) uint8, bit (m % 8) of byte (m / 8) for Morton code m
    (volume_func.h:183-194)."""
    n = vol.shape[0]
    m = morton_index(n).reshape(-1)
    flat = np.zeros(n ** 3, np.uint8)
    flat[m] = vol.reshape(-1).astype(np.uint8)
    return np.packbits(flat.reshape(-1, 8), axis=1, bitorder='little').reshape(-1)


def bitfield_to_volume(bits, n):
    flat = np.unpackbits(np.asarray(bits, np.uint8).reshape(-1, 1), axis=1, bitorder='little').reshape(-1)
    return flat[morton_index(n).reshape(-1)].reshape(n, n, n).astype(bool)
