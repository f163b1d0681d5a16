# -*- coding: utf-8 -*-
"""Object bounds of the foreground model, same classes, method names, arguments and returns as
arcnerf/models/base_modules/obj_bound/{basic_bound,sphere_bound,volume_bound,bitfield_bound}.py and the factory
`build_obj_bound` (obj_bound/__init__.py:25-62).  The bounds own the occupancy state (bool volume + opacity field, or
density grid + Morton bitfield), pick near/far per ray, produce the depth samples through libarcb200's march kernels and
refresh the occupancy in `optimize()` from a caller-supplied opacity estimator.

Configs may be the reference's nested `Obj` nodes or plain dicts.
"""
from copy import deepcopy

import torch
import torch.nn as nn

from .bitfield_func import (
    count_bitfield, ema_grid_samples_nerf, generate_grid_samples, sparse_volume_sampling_bit, splat_grid_samples,
    update_bitfield
)
from .ray_helper import (
    get_near_far_from_rays, get_zvals_from_near_far, get_zvals_from_near_far_fix_step, sphere_ray_intersection
)
from .registry import BOUND_REGISTRY
from .volume import Volume
from .volume_func import sparse_volume_sampling, tensor_reduce_max

CUDA_BACKEND_AVAILABLE = True


def _as_dict(cfgs):
    if cfgs is None:
        return {}
    return cfgs if isinstance(cfgs, dict) else cfgs.__dict__


def valid_key_in_cfgs(cfgs, key):
    """common/utils/cfgs_utils.py: the key exists (and the node is not None)"""
    return cfgs is not None and key in _as_dict(cfgs)


def get_value_from_cfgs_field(cfgs, key, default=None):
    return _as_dict(cfgs).get(key, default) if cfgs is not None else default


@BOUND_REGISTRY.register()
class BasicBound(nn.Module):
    """basic_bound.py:12-106: no structure; near/far from bounds / hard-coded values / a bounding sphere"""

    def __init__(self, cfgs):
        super(BasicBound, self).__init__()
        self.cfgs = cfgs
        self.optim_cfgs = self.read_optim_cfgs()

    def get_obj_bound(self):
        return None

    def get_optim_cfgs(self, key=None):
        return self.optim_cfgs if key is None else self.optim_cfgs[key]

    def set_optim_cfgs(self, key, value):
        self.optim_cfgs[key] = value

    def read_optim_cfgs(self):
        return {
            'epoch_optim': get_value_from_cfgs_field(self.cfgs, 'epoch_optim', None),
            'epoch_optim_warmup': get_value_from_cfgs_field(self.cfgs, 'epoch_optim_warmup', None),
            'ema_optim_decay': get_value_from_cfgs_field(self.cfgs, 'ema_optim_decay', 0.95),
            'opa_thres': get_value_from_cfgs_field(self.cfgs, 'opa_thres', 0.01)
        }

    def get_near_far_from_rays(self, inputs, near_hardcode=None, far_hardcode=None, bounding_radius=None):
        """-> near, far (B, 1), mask_rays None"""
        bounds = inputs['bounds'] if 'bounds' in inputs else None
        near, far = get_near_far_from_rays(inputs['rays_o'], inputs['rays_d'], bounds, near_hardcode, far_hardcode,
                                           bounding_radius)
        return near, far, None

    def get_zvals_from_near_far(self, near, far, n_pts, inference_only=False, inverse_linear=False, perturb=False, **kwargs):
        """-> zvals (B, n_pts), mask_pts None"""
        zvals = get_zvals_from_near_far(near, far, n_pts, inverse_linear=inverse_linear,
                                        perturb=perturb if not inference_only else False)
        return zvals, None

    @torch.no_grad()
    def optimize(self, cur_epoch=0, n_pts=128, get_est_opacity=None):
        return


class Sphere(nn.Module):
    """arcnerf/geometry/sphere.py, the part SphereBound needs: centre + radius and the ray test"""

    def __init__(self, origin=(0, 0, 0), radius=1.0, dtype=torch.float32, requires_grad=False, **kwargs):
        super(Sphere, self).__init__()
        self.origin = nn.Parameter(torch.tensor([float(x) for x in origin], dtype=dtype), requires_grad=requires_grad)
        self.radius = nn.Parameter(torch.tensor([float(radius)], dtype=dtype), requires_grad=requires_grad)

    def get_origin(self):
        return self.origin

    def get_radius(self, in_float=False):
        return float(self.radius[0]) if in_float else self.radius

    def ray_sphere_intersection(self, rays_o, rays_d):
        """-> near, far (N, 1), pts (N, 2, 3), mask (N, 1)"""
        near, far, pts, mask = sphere_ray_intersection(rays_o, rays_d, self.radius.detach(),
                                                       tuple(float(x) for x in self.origin.detach().cpu()))
        return near, far, pts[:, 0, :, :], mask


@BOUND_REGISTRY.register()
class SphereBound(BasicBound):
    """sphere_bound.py:9-37"""

    def __init__(self, cfgs):
        super(SphereBound, self).__init__(cfgs)
        assert valid_key_in_cfgs(cfgs, 'sphere'), 'You must have sphere in the cfgs'
        self.sphere = Sphere(**_as_dict(_as_dict(cfgs)['sphere']))

    def get_obj_bound(self):
        return self.sphere

    def get_near_far_from_rays(self, inputs, **kwargs):
        near, far, _, mask_rays = self.sphere.ray_sphere_intersection(inputs['rays_o'], inputs['rays_d'])
        return near, far, mask_rays[:, 0]


@BOUND_REGISTRY.register()
class VolumeBound(BasicBound):
    """volume_bound.py:15-212: bool occupancy volume + opacity field, pruned in optimize()"""

    def __init__(self, cfgs):
        super(VolumeBound, self).__init__(cfgs)
        assert valid_key_in_cfgs(cfgs, 'volume'), 'You must have volume in the cfgs'
        volume_cfgs = dict(_as_dict(_as_dict(cfgs)['volume']))
        if volume_cfgs.get('n_grid') is None:
            volume_cfgs['n_grid'] = 128
        self.volume = Volume(**volume_cfgs)
        if self.get_optim_cfgs('epoch_optim') is not None:
            self.volume.set_up_voxel_bitfield(init_occ=True)
            self.volume.set_up_voxel_opafield()

    def get_obj_bound(self):
        return self.volume

    def read_optim_cfgs(self):
        params = super().read_optim_cfgs()
        params['ray_sample_acc'] = get_value_from_cfgs_field(self.cfgs, 'ray_sample_acc', False)
        params['ray_sample_fix_step'] = get_value_from_cfgs_field(self.cfgs, 'ray_sample_fix_step', False)
        params['near_distance'] = get_value_from_cfgs_field(self.cfgs, 'near_distance', 0.0)
        return params

    def get_near_far_from_rays(self, inputs, **kwargs):
        near, far, _, mask_rays = self.volume.ray_volume_intersection(inputs['rays_o'], inputs['rays_d'])
        return near, far, mask_rays[:, 0]

    def get_zvals_from_near_far(self, near, far, n_pts, inference_only=False, inverse_linear=False, perturb=False,
                                rays_o=None, rays_d=None, **kwargs):
        if self.get_optim_cfgs('epoch_optim') is not None and self.get_optim_cfgs('ray_sample_acc'):
            return self.get_zvals_from_sparse_volume(rays_o, rays_d, near, far, n_pts, inference_only, inverse_linear, perturb)
        return super().get_zvals_from_near_far(near, far, n_pts, inference_only, inverse_linear, perturb)

    @torch.no_grad()
    def get_zvals_from_sparse_volume(self, rays_o, rays_d, near, far, n_pts, inference_only, inverse_linear, perturb):
        """-> zvals (B, n_pts), mask_pts (B, n_pts) bool: the fixed-step march that skips empty voxels"""
        const_dt = self.volume.get_diag_len() / n_pts
        return sparse_volume_sampling(rays_o, rays_d, near, far, n_pts, const_dt, self.volume.get_range(),
                                      self.volume.get_n_grid(), self.volume.get_voxel_bitfield().contiguous(),
                                      near_distance=self.get_optim_cfgs('near_distance'))

    def get_zvals_from_near_far_fix_step(self, near, far, n_pts, inference_only=False, perturb=False, **kwargs):
        fix_t = self.volume.get_diag_len() / n_pts
        return get_zvals_from_near_far_fix_step(near, far, fix_t, n_pts, perturb=perturb if not inference_only else False)

    @torch.no_grad()
    def optimize(self, cur_epoch=0, n_pts=128, get_est_opacity=None):
        """volume_bound.py:160-212.  Warm-up: every voxel; afterwards 1/4 random voxels + the first 1/4 occupied ones.
        One jittered point per chosen voxel -> opacity -> per-voxel max -> decayed-max update of the opacity field ->
        occupancy = opacity >= min(mean, threshold).
        Default: the device path (csrc/volume_refresh.cu: no randperm / where / unique / host round trip).  With
        `native_refresh: False` in the cfgs the selection and update are composed from torch ops in the reference's order."""
        epoch_optim = self.get_optim_cfgs('epoch_optim')
        epoch_optim_warmup = self.get_optim_cfgs('epoch_optim_warmup')
        if cur_epoch <= 0 or epoch_optim is None or cur_epoch % epoch_optim != 0:
            return
        vol = self.volume
        if get_value_from_cfgs_field(self.cfgs, 'native_refresh', True) and vol.get_range().is_cuda:
            warm = epoch_optim_warmup is not None and cur_epoch < epoch_optim_warmup
            n_vox = vol.get_n_voxel()
            return self._optimize_native(n_vox if warm else n_vox // 4, 0 if warm else n_vox // 4, n_pts, get_est_opacity)
        n_grid, dev = vol.get_n_grid(), vol.get_device()
        if epoch_optim_warmup is not None and cur_epoch < epoch_optim_warmup:
            flat_idx = torch.arange(vol.get_n_voxel(), dtype=torch.long, device=dev)
        else:
            n_sample = vol.get_n_voxel() // 4
            uni = torch.randperm(vol.get_n_voxel(), dtype=torch.long, device=dev)[:n_sample]
            occ = vol.get_occupied_voxel_idx(flatten=True)[:n_sample]
            flat_idx = torch.cat([uni, occ], dim=0)
        voxel_idx = vol.convert_flatten_index_to_xyz_index(flat_idx, n_grid)
        voxel_pts = vol.get_voxel_pts_by_voxel_idx(voxel_idx)
        noise = torch.rand_like(voxel_pts) - 0.5
        voxel_pts = voxel_pts + noise * vol.get_voxel_size(to_list=False)[None, :]
        dt = vol.get_diag_len() / float(n_pts)
        opacity = get_est_opacity(dt, voxel_pts)
        # max opacity per distinct voxel (sorted by flat index = the reference's lexicographic unique over (x, y, z))
        uni_flat, inverse = torch.unique(flat_idx, return_inverse=True)
        uni_opa = tensor_reduce_max(opacity.float(), inverse, uni_flat.shape[0])
        vol.update_opafield_by_voxel_idx(vol.convert_flatten_index_to_xyz_index(uni_flat, n_grid), uni_opa,
                                         ema=self.get_optim_cfgs('ema_optim_decay'))
        vol.update_bitfield_by_opafield(threshold=self.get_optim_cfgs('opa_thres'), ops='overwrite')


    def _optimize_native(self, n_uniform, n_occ, n_pts, get_est_opacity):
        from .. import _lib
        vol = self.volume
        n_grid, dev = vol.get_n_grid(), vol.get_device()
        n_cand = n_uniform + n_occ
        if getattr(self, '_refresh_scratch', None) is None or self._refresh_scratch.device != dev:
            nbytes = _lib.load().arcb200_volume_refresh_scratch_bytes(int(n_grid))
            self._refresh_scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            self._refresh_rng = _lib.Rng()
            self._refresh_count = 0
        if getattr(self, '_refresh_idx', None) is None or self._refresh_idx.shape[0] < n_cand:
            self._refresh_idx = torch.empty((n_cand, ), dtype=torch.int32, device=dev)
            self._refresh_pts = torch.empty((n_cand, 3), dtype=torch.float32, device=dev)
        idx, pts = self._refresh_idx[:n_cand], self._refresh_pts[:n_cand]
        bits = vol.get_voxel_bitfield()
        if not bits.is_contiguous():
            bits = bits.contiguous()
        if getattr(self, '_refresh_aabb', None) is None or self._refresh_aabb.device != dev:
            self._refresh_aabb = vol.get_range().t().contiguous()      # (2, 3): min; max
        aabb = self._refresh_aabb
        st = _lib.current_stream()
        if n_uniform == vol.get_n_voxel() and n_occ == 0:
            key = 0xffffffff                                           # warm-up: every voxel once (any permutation will do)
        else:
            key = (self._refresh_count * 2654435761) & 0xffffffff
        r = self._refresh_rng
        _lib.call('arcb200_volume_refresh_candidates', bits.data_ptr(), int(n_grid), n_uniform, n_occ, key, r.state, r.inc,
                  aabb.data_ptr(), self._refresh_scratch.data_ptr(), idx.data_ptr(), pts.data_ptr(), st)
        r.advance()
        self._refresh_count += 1
        dt = vol.get_diag_len() / float(n_pts)
        opacity = get_est_opacity(dt, pts).float().contiguous()
        ema = self.get_optim_cfgs('ema_optim_decay')
        opa = vol.get_voxel_opafield()
        _lib.call('arcb200_volume_refresh_apply', idx.data_ptr(), opacity.data_ptr(), n_cand, -1.0 if ema is None else float(ema),
                  float(self.get_optim_cfgs('opa_thres')), int(n_grid), self._refresh_scratch.data_ptr(), opa.data_ptr(),
                  bits.data_ptr(), st)
        if bits.data_ptr() != vol.bitfield.data_ptr():
            vol.bitfield.copy_(bits)


@BOUND_REGISTRY.register()
class BitfieldBound(BasicBound):
    """bitfield_bound.py:16-189: density grid + Morton bitfield, the instant-ngp pruning (one cascade)"""

    def __init__(self, cfgs):
        super(BitfieldBound, self).__init__(cfgs)
        assert valid_key_in_cfgs(cfgs, 'bitfield'), 'You must have bitfield in the cfgs'
        bitfield_cfgs = dict(_as_dict(_as_dict(cfgs)['bitfield']))
        if bitfield_cfgs.get('n_grid') is None:
            bitfield_cfgs['n_grid'] = 128
        self.volume = Volume(**bitfield_cfgs)
        self.n_grid = self.volume.get_n_grid()
        self.n_elements = self.n_grid ** 3
        if self.get_optim_cfgs('epoch_optim') is not None:
            self.register_buffer('density_bitfield', torch.full((self.n_elements // 8, ), 255, dtype=torch.uint8))
            self.register_buffer('density_grid', torch.zeros((self.n_elements, ), dtype=torch.float32))
            self.register_buffer('density_grid_tmp', torch.zeros((self.n_elements, ), dtype=torch.float32))
            self.ema_step = 0

    def get_obj_bound(self):
        return self.density_bitfield if self.get_optim_cfgs('epoch_optim') is not None else None

    def get_n_grid(self):
        return self.n_grid

    def read_optim_cfgs(self):
        params = super().read_optim_cfgs()
        params['near_distance'] = get_value_from_cfgs_field(self.cfgs, 'near_distance', 0.0)
        return params

    def get_near_far_from_rays(self, inputs, **kwargs):
        near, far, _, mask_rays = self.volume.ray_volume_intersection(inputs['rays_o'], inputs['rays_d'])
        return near, far, mask_rays[:, 0]

    def get_zvals_from_near_far(self, near, far, n_pts, inference_only=False, inverse_linear=False, perturb=False,
                                rays_o=None, rays_d=None, **kwargs):
        const_dt = self.volume.get_diag_len() / n_pts
        return sparse_volume_sampling_bit(rays_o, rays_d, near, far, n_pts, const_dt, self.volume.get_range(),
                                          self.volume.get_n_grid(), self.density_bitfield,
                                          near_distance=self.get_optim_cfgs('near_distance'))

    def get_density_grid_mean(self):
        """(1,) tensor holding mean(max(density, 0)); stays on the device (the reference reads it back with .item())"""
        return self.density_grid.clamp_min(0.0).mean().reshape(1)

    def get_bitfield_count(self, level=0):
        bitcount = count_bitfield(self.density_bitfield, self.n_grid)
        return bitcount, float(bitcount) / float(self.n_elements)

    @torch.no_grad()
    def optimize(self, cur_epoch=0, n_pts=128, get_est_opacity=None):
        epoch_optim = self.get_optim_cfgs('epoch_optim')
        epoch_optim_warmup = self.get_optim_cfgs('epoch_optim_warmup')
        if cur_epoch <= 0 or epoch_optim is None or cur_epoch % epoch_optim != 0:
            return
        if epoch_optim_warmup is not None and cur_epoch < epoch_optim_warmup:
            self._update_density_grid(self.n_elements, 0, get_est_opacity, n_pts)
        else:
            self._update_density_grid(self.n_elements // 4, self.n_elements // 4, get_est_opacity, n_pts)

    def _update_density_grid(self, n_uniform, n_nonuniform, get_est_opacity, n_pts):
        """bitfield_bound.py:148-189"""
        n_total_sample = n_uniform + n_nonuniform
        pos_uniform, idx_uniform = generate_grid_samples(self.density_grid, n_uniform, self.ema_step, self.n_grid, -0.01)
        pos_nonuniform, idx_nonuniform = generate_grid_samples(self.density_grid, n_nonuniform, self.ema_step, self.n_grid,
                                                               self.get_optim_cfgs('opa_thres'))
        pos_sample = torch.cat([pos_uniform, pos_nonuniform], dim=0)
        idx_sample = torch.cat([idx_uniform, idx_nonuniform], dim=0)
        rng = self.volume.get_range()
        pos_sample = pos_sample * (rng[:, 1] - rng[:, 0])[None] + rng[:, 0][None]
        dt = self.volume.get_diag_len() / float(n_pts)
        opacity = get_est_opacity(dt, pos_sample)
        self.density_grid_tmp.zero_()
        self.density_grid_tmp = splat_grid_samples(opacity, idx_sample, n_total_sample, self.density_grid_tmp)
        self.density_grid = ema_grid_samples_nerf(self.density_grid_tmp, self.density_grid, self.n_elements,
                                                  self.get_optim_cfgs('ema_optim_decay'))
        self.density_bitfield = update_bitfield(self.density_grid, self.get_density_grid_mean(), self.density_bitfield,
                                                self.get_optim_cfgs('opa_thres'), self.n_grid)
        self.ema_step = self.ema_step + 1


def key_to_bound_type(key):
    return {'basic': 'BasicBound', 'bitfield': 'BitfieldBound', 'volume': 'VolumeBound', 'sphere': 'SphereBound'}[key]


def build_obj_bound(cfgs):
    """obj_bound/__init__.py:25-51: -> (bound module, its type key) from cfgs.obj_bound"""
    if not valid_key_in_cfgs(cfgs, 'obj_bound'):
        key = 'basic'
        obj_bound = BOUND_REGISTRY.get(key_to_bound_type(key))(None)
    else:
        keys = _as_dict(get_value_from_cfgs_field(cfgs, 'obj_bound')).keys()
        if 'volume' in keys:
            key = 'volume'
        elif 'sphere' in keys:
            key = 'sphere'
        elif 'bitfield' in keys:
            key = 'bitfield'
        else:
            raise NotImplementedError('Not such bounding class {}...'.format(keys))
        obj_bound = BOUND_REGISTRY.get(key_to_bound_type(key))(deepcopy(_as_dict(cfgs)['obj_bound']))
    return obj_bound, key
