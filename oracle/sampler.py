"""ctypes front-end of the C sampler oracle (oracle/sampler_oracle.c) and, when it has been built in
a container that holds /root/reference, of the host build of the reference's own sampler functions
(oracle/_ref/libarcref.so, see oracle/Makefile).  TEST INFRASTRUCTURE ONLY.

Both libraries expose the same call shapes; ``Sampler('oracle')`` / ``Sampler('ref')`` pick one.
Arrays are numpy, C-contiguous; aabb is (2,3) [min xyz; max xyz] as the reference's C++ layer sees it
(the python wrappers of the reference permute their (3,2) range before the call,
arcnerf/ops/volume_func.py:54-56,95-98).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ORACLE_SO = os.path.join(_HERE, '_build', 'liboracle.so')
_REF_SO = os.path.join(_HERE, '_ref', 'libarcref.so')

RNG_SEED = 9121  # arcnerf/ops/include/common.h:23


def build(ref=True):
    """Compile the C oracle (always) and the reference host build (when /root/reference exists)."""
    subprocess.check_call(['make', '-s', '-C', _HERE, 'oracle'])
    if ref and os.path.isdir('/root/reference/arcnerf/ops/include'):
        subprocess.check_call(['make', '-s', '-C', _HERE, 'ref'])
        # the reference's device kernels for sm_100 (run only on the GPU box, by scripts/ref_kernels_on_box.py)
        subprocess.check_call(['make', '-s', '-C', _HERE, 'ref_dev'])
        # tiny-cuda-nn for sm_100 takes about 15 minutes from scratch: only on request (the golden vectors it produced are
        # committed, tests never need the library itself)
        if os.environ.get('ARCB200_BUILD_TCNN') == '1':
            subprocess.check_call(['make', '-s', '-j4', '-C', _HERE, 'tcnn'])


def have_ref():
    return os.path.exists(_REF_SO)


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float32)


class Rng:
    """The per-module generator of the reference (`static pcg32 rng{9121}`), advanced by 2**32 after
    every launch that consumes it (volume_func_kernel.cu:288, bitfield_func_kernel.cu:134,209)."""

    def __init__(self, lib, prefix, seed=RNG_SEED):
        self._lib, self._pre = lib, prefix
        s, i = C.c_uint64(), C.c_uint64()
        getattr(lib, prefix + 'rng_init')(C.c_uint64(seed), C.byref(s), C.byref(i))
        self.state, self.inc = s.value, i.value

    def advance(self, delta=1 << 32):
        s = C.c_uint64(self.state)
        getattr(self._lib, self._pre + 'rng_advance')(C.byref(s), C.c_uint64(self.inc), C.c_int64(delta))
        self.state = s.value

    def next_float(self):
        s = C.c_uint64(self.state)
        fn = getattr(self._lib, self._pre + 'rng_next_float')
        fn.restype = C.c_float
        v = fn(C.byref(s), C.c_uint64(self.inc))
        self.state = s.value
        return float(v)


class Sampler:
    def __init__(self, which='oracle'):
        if which == 'oracle':
            if not os.path.exists(_ORACLE_SO):
                build(ref=False)
            self.lib, self.pre = C.CDLL(_ORACLE_SO), 'orc_'
        elif which == 'ref':
            if not os.path.exists(_REF_SO):
                raise FileNotFoundError('oracle/_ref/libarcref.so not built (needs /root/reference)')
            self.lib, self.pre = C.CDLL(_REF_SO), 'ref_'
        else:
            raise ValueError(which)
        self.which = which

    def rng(self, seed=RNG_SEED):
        return Rng(self.lib, self.pre, seed)

    def _fn(self, name):
        return getattr(self.lib, self.pre + name)

    def morton3D(self, x, y, z):
        fn = self._fn('morton3D'); fn.restype = C.c_uint32
        return int(fn(C.c_uint32(x), C.c_uint32(y), C.c_uint32(z)))

    def morton3D_invert(self, x):
        fn = self._fn('morton3D_invert'); fn.restype = C.c_uint32
        return int(fn(C.c_uint32(x)))

    def aabb_intersection(self, rays_o, rays_d, aabb):
        """aabb: (N_v, 2, 3).  Returns near, far (N, N_v), pts (N, N_v, 2, 3), mask (N, N_v) bool."""
        o, d, ab = _f(rays_o), _f(rays_d), _f(aabb)
        n, nv = o.shape[0], ab.shape[0]
        near = np.zeros((n, nv), np.float32); far = np.zeros((n, nv), np.float32)
        pts = np.zeros((n, nv, 2, 3), np.float32); mask = np.zeros((n, nv), np.uint8)
        self._fn('aabb_intersection')(n, nv, _p(o, C.c_float), _p(d, C.c_float), _p(ab, C.c_float),
                                      _p(near, C.c_float), _p(far, C.c_float), _p(pts, C.c_float),
                                      _p(mask, C.c_uint8))
        return near, far, pts, mask.astype(bool)

    def check_pts_in_occ_voxel(self, xyz, volume, aabb, n_grid):
        x, ab = _f(xyz), _f(aabb)
        vol = np.ascontiguousarray(volume, dtype=np.uint8)
        out = np.zeros((x.shape[0],), np.uint8)
        self._fn('check_pts_in_occ_voxel')(x.shape[0], _p(x, C.c_float), _p(vol, C.c_uint8),
                                           _p(ab, C.c_float), int(n_grid), _p(out, C.c_uint8))
        return out.astype(bool)

    def sparse_volume_sampling(self, rays_o, rays_d, near, far, n_pts, dt, aabb, n_grid, field,
                               near_distance, rng, bit=False):
        """field: (n,n,n) bool volume (bit=False) or (n**3/8,) uint8 Morton bitfield (bit=True).
        Does NOT advance `rng` (the launcher does that after the kernel)."""
        o, d, ab = _f(rays_o), _f(rays_d), _f(aabb)
        nr, fr = _f(near).reshape(-1), _f(far).reshape(-1)
        fld = np.ascontiguousarray(field, dtype=np.uint8)
        n = o.shape[0]
        z = np.zeros((n, n_pts), np.float32); m = np.zeros((n, n_pts), np.uint8)
        self._fn('sparse_volume_sampling')(
            int(bit), n, _p(o, C.c_float), _p(d, C.c_float), _p(nr, C.c_float), _p(fr, C.c_float),
            _p(ab, C.c_float), _p(fld, C.c_uint8), int(n_grid), int(n_pts), C.c_float(dt),
            C.c_float(near_distance), C.c_uint64(rng.state), C.c_uint64(rng.inc), _p(z, C.c_float),
            _p(m, C.c_uint8))
        return z, m.astype(bool)

    def sparse_sampling_in_multivol_bitfield(self, rays_o, rays_d, near, far, n_pts, cone_angle, min_step,
                                             max_step, min_aabb, aabb, n_grid, n_cascade, bitfield,
                                             near_distance, inclusive, rng):
        o, d, ab, mab = _f(rays_o), _f(rays_d), _f(aabb), _f(min_aabb)
        nr, fr = _f(near).reshape(-1), _f(far).reshape(-1)
        fld = np.ascontiguousarray(bitfield, dtype=np.uint8)
        n = o.shape[0]
        z = np.zeros((n, n_pts), np.float32); m = np.zeros((n, n_pts), np.uint8)
        self._fn('sparse_sampling_in_multivol_bitfield')(
            n, _p(o, C.c_float), _p(d, C.c_float), _p(nr, C.c_float), _p(fr, C.c_float),
            _p(mab, C.c_float), _p(ab, C.c_float), _p(fld, C.c_uint8), int(n_grid), int(n_cascade),
            int(n_pts), C.c_float(cone_angle), C.c_float(min_step), C.c_float(max_step),
            C.c_float(near_distance), int(bool(inclusive)), C.c_uint64(rng.state), C.c_uint64(rng.inc),
            _p(z, C.c_float), _p(m, C.c_uint8))
        return z, m.astype(bool)

    def generate_grid_samples(self, grid, step, n_elements, n_grid, thresh, rng):
        g = _f(grid)
        pos = np.zeros((n_elements, 3), np.float32); idx = np.zeros((n_elements,), np.uint32)
        self._fn('generate_grid_samples')(int(n_elements), C.c_uint64(rng.state), C.c_uint64(rng.inc),
                                          C.c_uint32(step), _p(g, C.c_float), _p(pos, C.c_float),
                                          _p(idx, C.c_uint32), C.c_uint32(n_grid), C.c_float(thresh))
        return pos, idx

    def generate_grid_samples_multivol(self, grid, step, n_elements, aabb, n_cascade, n_grid, thresh,
                                       inclusive, rng):
        g, ab = _f(grid), _f(aabb)
        pos = np.zeros((n_elements, 3), np.float32); idx = np.zeros((n_elements,), np.uint32)
        self._fn('generate_grid_samples_multivol')(
            int(n_elements), _p(ab, C.c_float), C.c_uint64(rng.state), C.c_uint64(rng.inc),
            C.c_uint32(step), _p(g, C.c_float), _p(pos, C.c_float), _p(idx, C.c_uint32),
            C.c_uint32(n_cascade), C.c_uint32(n_grid), C.c_float(thresh), int(bool(inclusive)))
        return pos, idx

    # ---- oracle-only helpers (the reference bodies are trivially elementwise) ----
    def sparse_volume_count(self, rays_o, rays_d, near, far, n_pts, dt, aabb, n_grid, field,
                            near_distance, rng, bit=False):
        assert self.which == 'oracle'
        o, d, ab = _f(rays_o), _f(rays_d), _f(aabb)
        nr, fr = _f(near).reshape(-1), _f(far).reshape(-1)
        fld = np.ascontiguousarray(field, dtype=np.uint8)
        n = o.shape[0]
        cnt = np.zeros((n,), np.int32)
        self.lib.orc_sparse_volume_count(
            int(bit), n, _p(o, C.c_float), _p(d, C.c_float), _p(nr, C.c_float), _p(fr, C.c_float),
            _p(ab, C.c_float), _p(fld, C.c_uint8), int(n_grid), int(n_pts), C.c_float(dt),
            C.c_float(near_distance), C.c_uint64(rng.state), C.c_uint64(rng.inc), _p(cnt, C.c_int32))
        return cnt

    def tensor_reduce_max(self, full, group, n_group):
        assert self.which == 'oracle'
        f = _f(full); g = np.ascontiguousarray(group, dtype=np.int64)
        uni = np.zeros((n_group,), np.float32)
        self.lib.orc_tensor_reduce_max(f.shape[0], _p(f, C.c_float), _p(g, C.c_int64), int(n_group),
                                       _p(uni, C.c_float))
        return uni

    def splat_grid_samples(self, density, indices, grid_tmp):
        assert self.which == 'oracle'
        dn = _f(density); ix = np.ascontiguousarray(indices, dtype=np.uint32)
        out = _f(grid_tmp).copy()
        self.lib.orc_splat_grid_samples(dn.shape[0], _p(ix, C.c_uint32), _p(dn, C.c_float), _p(out, C.c_float))
        return out

    def ema_grid_samples_nerf(self, grid_tmp, decay, grid):
        assert self.which == 'oracle'
        t = _f(grid_tmp); g = _f(grid).copy()
        self.lib.orc_ema_grid_samples_nerf(g.shape[0], C.c_float(decay), _p(g, C.c_float), _p(t, C.c_float))
        return g

    def grid_to_bitfield(self, grid, mean_density, opa_thres):
        assert self.which == 'oracle'
        g = _f(grid); bits = np.zeros((g.shape[0] // 8,), np.uint8)
        self.lib.orc_grid_to_bitfield(bits.shape[0], _p(g, C.c_float), _p(bits, C.c_uint8),
                                      C.c_float(mean_density), C.c_float(opa_thres))
        return bits

    def count_bitfield(self, bits, faithful=True):
        assert self.which == 'oracle'
        b = np.ascontiguousarray(bits, dtype=np.uint8); c = np.zeros((1,), np.float32)
        self.lib.orc_count_bitfield(b.shape[0], _p(b, C.c_uint8), _p(c, C.c_float), int(faithful))
        return float(c[0])
