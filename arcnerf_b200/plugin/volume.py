# -*- coding: utf-8 -*-
"""The part of arcnerf/geometry/volume.py:Volume that the object bounds use: box geometry, the bool
occupancy volume (`bitfield`), the per-voxel opacity (`opafield`) and their update rules, with the ray/box test and the
point-in-occupied-voxel test going to libarcb200 (arcnerf_b200/plugin/volume_func.py).

Same method names, arguments and return layouts as the reference; geometry helpers that only serve visualisation or
mesh extraction (corner/line/face lists, trilinear grid interpolation) are not part of the hot path and are left out.
"""
import torch
import torch.nn as nn

from .volume_func import check_pts_in_occ_voxel_cuda, ray_aabb_intersection_cuda


class Volume(nn.Module):
    """volume.py:15-50: n_grid**3 voxels in a box centred at `origin` with side length(s) `side` / `xyz_len`"""

    def __init__(self, n_grid=None, origin=(0, 0, 0), side=None, xyz_len=None, dtype=torch.float32, requires_grad=False,
                 **kwargs):
        super(Volume, self).__init__()
        self.requires_grad = requires_grad
        self.dtype = dtype
        self.n_grid = n_grid
        self.origin = nn.Parameter(torch.tensor([0.0, 0.0, 0.0], dtype=dtype), requires_grad=requires_grad)
        self.xyz_len = nn.Parameter(torch.tensor([0.0, 0.0, 0.0], dtype=dtype), requires_grad=requires_grad)
        self.contains_bitfield = False
        if origin is not None and (side is not None or xyz_len is not None):
            self.set_params(origin, side, xyz_len)

    # ------------------------------------------------------------------ geometry (volume.py:52-253)
    @torch.no_grad()
    def set_params(self, origin, side, xyz_len):
        assert side is not None or xyz_len is not None, 'Specify at least side or xyz_len'
        self.origin.copy_(torch.tensor([float(origin[0]), float(origin[1]), float(origin[2])], dtype=self.dtype))
        lens = [float(side)] * 3 if side is not None else [float(x) for x in xyz_len]
        self.xyz_len.copy_(torch.tensor(lens, dtype=self.dtype))
        # host copies of the box: the getters below are called every refresh and must not read device memory back
        self._host_min = [float(origin[a]) - lens[a] / 2.0 for a in range(3)]
        self._host_max = [float(origin[a]) + lens[a] / 2.0 for a in range(3)]
        self.set_pts()

    def set_pts(self):
        self._diag_len = None
        self.cal_range()
        if self.n_grid is not None:
            self.cal_volume_pts()

    def get_n_grid(self):
        return self.n_grid

    def get_n_voxel(self):
        return self.n_grid ** 3

    def get_device(self):
        return self.origin.device

    def get_len(self):
        return float(self.xyz_len[0]), float(self.xyz_len[1]), float(self.xyz_len[2])

    def get_origin(self):
        return self.origin

    def cal_range(self):
        xyz_min = self.origin.detach() - self.xyz_len.detach() / 2.0
        xyz_max = self.origin.detach() + self.xyz_len.detach() / 2.0
        self.register_buffer('range', torch.cat([xyz_min[:, None], xyz_max[:, None]], dim=-1))   # (3, 2)

    def get_range(self):
        return self.range

    def get_diag_len(self):
        """length of the box diagonal, evaluated like the reference (fp32 tensor arithmetic) but once, on the host"""
        if getattr(self, '_diag_len', None) is None:
            r = torch.tensor([self._host_min, self._host_max], dtype=self.dtype).t()
            self._diag_len = float(torch.sqrt(((r[:, 1] - r[:, 0]) ** 2).sum()))
        return self._diag_len

    def get_voxel_size(self, to_list=True):
        if to_list:
            r = torch.tensor([self._host_min, self._host_max], dtype=self.dtype).t()
            xyz_s = (r[:, 1] - r[:, 0]) / self.n_grid
            return float(xyz_s[0]), float(xyz_s[1]), float(xyz_s[2])
        return (self.get_range()[:, 1] - self.get_range()[:, 0]) / self.n_grid

    def cal_volume_pts(self):
        """centres of all voxels, (n_grid**3, 3), x-major flat order (x*n*n + y*n + z)"""
        v_x, v_y, v_z = self.get_voxel_size()
        n = self.n_grid
        dev = self.range.device
        x = torch.linspace(float(self.range[0, 0]) + 0.5 * v_x, float(self.range[0, 1]) - 0.5 * v_x, n, device=dev)
        y = torch.linspace(float(self.range[1, 0]) + 0.5 * v_y, float(self.range[1, 1]) - 0.5 * v_y, n, device=dev)
        z = torch.linspace(float(self.range[2, 0]) + 0.5 * v_z, float(self.range[2, 1]) - 0.5 * v_z, n, device=dev)
        volume_pts = torch.stack(torch.meshgrid(x, y, z, indexing='ij'), -1)
        self.register_buffer('volume_pts', volume_pts.view(-1, 3))

    def get_volume_pts(self, in_grid=False):
        if in_grid:
            return self.volume_pts.view(self.n_grid, self.n_grid, self.n_grid, 3)
        return self.volume_pts

    def get_full_voxel_idx(self, flatten=False):
        r = torch.arange(self.n_grid, dtype=torch.long, device=self.get_device())
        voxel_idx = torch.stack(torch.meshgrid(r, r, r, indexing='ij'), -1)
        return voxel_idx.view(-1, 3) if flatten else voxel_idx

    def get_voxel_pts_by_voxel_idx(self, voxel_idx):
        voxel_size = self.get_voxel_size(to_list=False)
        return voxel_idx * voxel_size + 0.5 * voxel_size + self.get_range()[:, 0]

    @staticmethod
    def convert_xyz_index_to_flatten_index(xyz, n):
        return xyz[:, 0] * (n ** 2) + xyz[:, 1] * n + xyz[:, 2]

    @staticmethod
    def convert_flatten_index_to_xyz_index(flatten_index, n):
        x = torch.div(flatten_index, n * n, rounding_mode='trunc')
        y = torch.div(flatten_index, n, rounding_mode='trunc') % n
        return torch.stack([x, y, flatten_index % n], dim=1)

    # ------------------------------------------------------------------ rays (volume.py:624-651)
    def ray_volume_intersection(self, rays_o, rays_d, in_occ_voxel=False, force=False):
        """-> near, far (N_rays, 1), pts (N_rays, 2, 3), mask (N_rays, 1) bool against the outer box"""
        if in_occ_voxel:
            raise NotImplementedError('per-voxel ray intersection is not on the hot path; sample with the sparse march instead')
        aabb_range = self.get_range()[None].to(rays_o.device)
        near, far, pts, mask = ray_aabb_intersection_cuda(rays_o, rays_d, aabb_range)
        return near, far, pts[:, 0, :, :], mask

    # ------------------------------------------------------------------ occupancy (volume.py:741-957)
    def set_up_voxel_bitfield(self, init_occ=True):
        assert self.n_grid is not None, 'Voxel bitfield must be set for known resolution volume'
        self.contains_bitfield = True
        n = self.n_grid
        fill = torch.ones if init_occ else torch.zeros
        self.register_buffer('bitfield', fill((n, n, n), dtype=torch.bool, device=self.get_device()))

    def get_voxel_bitfield(self, flatten=False):
        if not self.contains_bitfield:
            return None
        return self.bitfield.view(-1) if flatten else self.bitfield

    def reset_voxel_bitfield(self, occ=True):
        self.bitfield = torch.ones_like(self.bitfield) if occ else torch.zeros_like(self.bitfield)

    def update_bitfield(self, occupancy, ops='and'):
        if len(occupancy.shape) == 1:
            assert occupancy.shape[0] == self.get_n_voxel(), 'Occupancy should cover each voxel...'
            n = round(occupancy.shape[0] ** (1 / 3))
            occupancy = occupancy.view(n, n, n)
        assert occupancy.dtype == torch.bool, 'Must input bool tensor'
        if ops == 'and':
            self.bitfield = torch.logical_and(self.bitfield, occupancy)
        elif ops == 'or':
            self.bitfield = torch.logical_or(self.bitfield, occupancy)
        elif ops == 'overwrite':
            self.bitfield = occupancy
        else:
            raise NotImplementedError('Ops {} not support for update bitfield...'.format(ops))

    def get_n_occupied_voxel(self):
        return self.bitfield.sum()

    def get_occupied_voxel_idx(self, flatten=False):
        occ = torch.where(self.get_voxel_bitfield(flatten=True))[0]
        return occ if flatten else self.convert_flatten_index_to_xyz_index(occ, self.n_grid)

    def check_pts_in_occ_voxel(self, pts):
        """(B, 3) -> (B,) bool, by the library kernel (the reference's torch fallback builds a (B, N_occ, 3) tensor)"""
        return check_pts_in_occ_voxel_cuda(pts, self.bitfield.contiguous(), self.get_range(), self.n_grid)

    # ------------------------------------------------------------------ opacity (volume.py:959-1018)
    def set_up_voxel_opafield(self):
        assert self.n_grid is not None, 'Voxel bitfield must be set for known resolution volume'
        n = self.n_grid
        self.register_buffer('opafield', torch.zeros((n, n, n), dtype=self.dtype, device=self.get_device()))

    def get_voxel_opafield(self, flatten=False):
        return self.opafield.view(-1) if flatten else self.opafield

    def update_opafield_by_voxel_idx(self, voxel_idx, opacity, ema=None):
        cur = self.opafield[voxel_idx[:, 0], voxel_idx[:, 1], voxel_idx[:, 2]]
        update_opa = opacity if ema is None else torch.max(cur * ema, opacity)
        self.opafield[voxel_idx[:, 0], voxel_idx[:, 1], voxel_idx[:, 2]] = torch.where(cur >= 0, update_opa, cur)

    def get_mean_voxel_opacity(self):
        return float(self.opafield.clamp(min=0).mean())

    def get_mean_occ_voxel_opacity(self):
        return float(self.opafield[self.get_voxel_bitfield(flatten=False)].clamp(min=0).mean())

    def update_bitfield_by_opafield(self, threshold=0.01, ops='and'):
        thres = min(self.get_mean_voxel_opacity(), threshold)
        self.update_bitfield(self.opafield >= thres, ops)
