"""Generates tests/golden/sampler_golden.npz from the REFERENCE's own sampler functions, host-built from
/root/reference by oracle/Makefile (oracle/_ref/libarcref.so).  Run in the build container only:

    python tests/golden/make_golden_sampler.py

The committed .npz lets the oracle and the CUDA kernels be checked against reference outputs on machines
where /root/reference (and hence oracle/_ref) may be absent.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))

import cases  # noqa: E402
from arcnerf_b200 import synth  # noqa: E402
from oracle.sampler import Sampler, build  # noqa: E402


def main():
    build(ref=True)
    ref = Sampler('ref')
    out = {}
    # rng bookkeeping: first draws and a 2^32 jump of the module generator
    r = ref.rng()
    out['rng_state0'] = np.array([r.state, r.inc], np.uint64)
    out['rng_first'] = np.array([r.next_float() for _ in range(4)], np.float32)
    r = ref.rng(); r.advance()
    out['rng_state_after_launch'] = np.array([r.state], np.uint64)

    o, d = cases.aabb_table()
    aabb = np.stack([synth.aabb_of_side(2.0), np.array([[-0.5, -0.25, 0.0], [0.75, 1.0, 1.5]], np.float32)], 0)
    for k, v in zip(['near', 'far', 'pts', 'mask'], ref.aabb_intersection(o, d, aabb)):
        out['aabb_table_' + k] = v

    for name in cases.MARCH_GOLDEN_CASES:
        c = cases.march_case(name)
        near, far, _, _ = ref.aabb_intersection(c['o'], c['d'], c['aabb'][None])
        bits = synth.volume_to_bitfield(c['vol'])
        for bit, fld in ((0, c['vol']), (1, bits)):
            rng = ref.rng()
            if bit:
                rng.advance()   # second launch of a session: generator moved on by 2^32
            z, m = ref.sparse_volume_sampling(c['o'], c['d'], near, far, c['n_pts'], c['dt'], c['aabb'], c['n_grid'],
                                              fld, c['near_distance'], rng, bit=bool(bit))
            cnt = m.sum(1).astype(np.int32)
            out['march_%s_bit%d_counts' % (name, bit)] = cnt
            # keep fixtures small: valid samples only (row-major concatenation) + the padded value per ray
            out['march_%s_bit%d_valid' % (name, bit)] = z[m]
            out['march_%s_bit%d_lastcol' % (name, bit)] = z[:, -1].copy()
        pts = (c['o'][:, None] + c['d'][:, None] * np.linspace(0.5, 5.0, 24, dtype=np.float32)[None, :, None]).reshape(-1, 3)
        out['occ_pts_%s' % name] = ref.check_pts_in_occ_voxel(pts, c['vol'], c['aabb'], c['n_grid'])

    for inclusive in (False, True):
        c = cases.multivol_case(inclusive)
        near = np.zeros((c['o'].shape[0], 1), np.float32)
        far = np.full((c['o'].shape[0], 1), 12.0, np.float32)
        z, m = ref.sparse_sampling_in_multivol_bitfield(
            c['o'], c['d'], near, far, c['n_pts'], c['cone'], c['smin'], c['smax'], c['min_aabb'], c['aabb'],
            c['n_grid'], c['n_cascade'], c['bits'], c['near_distance'], inclusive, ref.rng())
        out['multivol_incl%d_z' % inclusive] = z
        out['multivol_incl%d_m' % inclusive] = m

    rng = np.random.RandomState(31)
    grid = rng.uniform(-0.01, 0.05, size=(16 ** 3,)).astype(np.float32)
    pos, idx = ref.generate_grid_samples(grid, 3, 1000, 16, 0.01, ref.rng())
    out['gridsamples_pos'], out['gridsamples_idx'] = pos, idx
    gridm = rng.uniform(-0.01, 0.05, size=(16 ** 3 * 3,)).astype(np.float32)
    pos, idx = ref.generate_grid_samples_multivol(gridm, 5, 1000, synth.aabb_of_side(2.0), 4, 16, 0.01, False, ref.rng())
    out['gridsamples_mv_pos'], out['gridsamples_mv_idx'] = pos, idx
    out['morton'] = np.array([ref.morton3D(x, y, z) for x, y, z in [(0, 0, 0), (1, 2, 3), (127, 5, 99), (31, 31, 31)]], np.uint32)

    path = os.path.join(ROOT, 'tests', 'golden', 'sampler_golden.npz')
    np.savez_compressed(path, **out)
    print('wrote', path, os.path.getsize(path), 'bytes')


if __name__ == '__main__':
    main()
