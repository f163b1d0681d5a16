"""CPU: the C sampler oracle against (a) golden vectors generated from the reference's own host-built
sampler functions (tests/golden/sampler_golden.npz) and (b) that host build itself when it is present
(oracle/_ref, only where /root/reference exists).  Bit-exact everywhere."""
import os

import numpy as np
import pytest

import cases
from arcnerf_b200 import synth
from oracle.sampler import Sampler, have_ref

GOLD = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'sampler_golden.npz'))


@pytest.fixture(scope='module')
def orc():
    return Sampler('oracle')


def test_rng_stream(orc):
    r = orc.rng()
    assert np.array_equal(np.array([r.state, r.inc], np.uint64), GOLD['rng_state0'])
    assert np.array_equal(np.array([r.next_float() for _ in range(4)], np.float32), GOLD['rng_first'])
    r = orc.rng(); r.advance()
    assert r.state == int(GOLD['rng_state_after_launch'][0])


def test_morton(orc):
    got = np.array([orc.morton3D(x, y, z) for x, y, z in [(0, 0, 0), (1, 2, 3), (127, 5, 99), (31, 31, 31)]], np.uint32)
    assert np.array_equal(got, GOLD['morton'])
    for m in (0, 1, 12345, 2 ** 21 - 1):
        x, y, z = orc.morton3D_invert(m), orc.morton3D_invert(m >> 1), orc.morton3D_invert(m >> 2)
        assert orc.morton3D(x, y, z) == m
    # synth's numpy Morton table agrees
    tab = synth.morton_index(8)
    assert tab[3, 5, 6] == orc.morton3D(3, 5, 6)


def test_aabb_table(orc):
    o, d = cases.aabb_table()
    aabb = np.stack([synth.aabb_of_side(2.0), np.array([[-0.5, -0.25, 0.0], [0.75, 1.0, 1.5]], np.float32)], 0)
    near, far, pts, mask = orc.aabb_intersection(o, d, aabb)
    assert np.array_equal(mask, GOLD['aabb_table_mask'])
    assert np.array_equal(near, GOLD['aabb_table_near']) and np.array_equal(far, GOLD['aabb_table_far'])
    assert np.array_equal(pts, GOLD['aabb_table_pts'], equal_nan=True)
    # origin-inside rays are misses in the CUDA flavour of the reference (volume_func_kernel.cu:106-114)
    assert not mask[0, 0] and not mask[1, 0]


@pytest.mark.parametrize('name', cases.MARCH_GOLDEN_CASES)
def test_march_golden(orc, name):
    c = cases.march_case(name)
    near, far, _, _ = orc.aabb_intersection(c['o'], c['d'], c['aabb'][None])
    bits = synth.volume_to_bitfield(c['vol'])
    for bit, fld in ((0, c['vol']), (1, bits)):
        rng = orc.rng()
        if bit:
            rng.advance()
        z, m = orc.sparse_volume_sampling(c['o'], c['d'], near, far, c['n_pts'], c['dt'], c['aabb'], c['n_grid'], fld,
                                          c['near_distance'], rng, bit=bool(bit))
        cnt = m.sum(1).astype(np.int32)
        assert np.array_equal(cnt, GOLD['march_%s_bit%d_counts' % (name, bit)])
        assert np.array_equal(z[m], GOLD['march_%s_bit%d_valid' % (name, bit)])
        assert np.array_equal(z[:, -1], GOLD['march_%s_bit%d_lastcol' % (name, bit)])
        # structure: valid samples form a prefix, strictly increasing; the tail repeats the last one
        for i in range(z.shape[0]):
            k = cnt[i]
            assert m[i, :k].all() and not m[i, k:].any()
            assert np.all(np.diff(z[i, :k]) > 0)
            if 0 < k < c['n_pts']:
                assert np.all(z[i, k:] == z[i, k - 1])
            if k == 0:
                assert np.all(z[i] == 0)
        cnt2 = orc.sparse_volume_count(c['o'], c['d'], near, far, c['n_pts'], c['dt'], c['aabb'], c['n_grid'], fld,
                                       c['near_distance'], rng, bit=bool(bit))
        assert np.array_equal(cnt, cnt2)
    if name == 'full_cap':
        assert cnt.max() == c['n_pts']
    if name == 'empty':
        assert cnt.max() == 0
    pts = (c['o'][:, None] + c['d'][:, None] * np.linspace(0.5, 5.0, 24, dtype=np.float32)[None, :, None]).reshape(-1, 3)
    assert np.array_equal(orc.check_pts_in_occ_voxel(pts, c['vol'], c['aabb'], c['n_grid']), GOLD['occ_pts_%s' % name])


@pytest.mark.parametrize('inclusive', [False, True])
def test_multivol_golden(orc, inclusive):
    c = cases.multivol_case(inclusive)
    near = np.zeros((c['o'].shape[0], 1), np.float32)
    far = np.full((c['o'].shape[0], 1), 12.0, np.float32)
    z, m = orc.sparse_sampling_in_multivol_bitfield(
        c['o'], c['d'], near, far, c['n_pts'], c['cone'], c['smin'], c['smax'], c['min_aabb'], c['aabb'], c['n_grid'],
        c['n_cascade'], c['bits'], c['near_distance'], inclusive, orc.rng())
    assert np.array_equal(m, GOLD['multivol_incl%d_m' % inclusive])
    assert np.array_equal(z, GOLD['multivol_incl%d_z' % inclusive])
    assert m.sum() > 0


def test_grid_samples_golden(orc):
    rng = np.random.RandomState(31)
    grid = rng.uniform(-0.01, 0.05, size=(16 ** 3,)).astype(np.float32)
    pos, idx = orc.generate_grid_samples(grid, 3, 1000, 16, 0.01, orc.rng())
    assert np.array_equal(idx, GOLD['gridsamples_idx']) and np.array_equal(pos, GOLD['gridsamples_pos'])
    gridm = rng.uniform(-0.01, 0.05, size=(16 ** 3 * 3,)).astype(np.float32)
    pos, idx = orc.generate_grid_samples_multivol(gridm, 5, 1000, synth.aabb_of_side(2.0), 4, 16, 0.01, False, orc.rng())
    assert np.array_equal(idx, GOLD['gridsamples_mv_idx']) and np.array_equal(pos, GOLD['gridsamples_mv_pos'])


def test_grid_maintenance_properties(orc):
    rng = np.random.RandomState(4)
    n = 16 ** 3
    grid = rng.uniform(-0.5, 1.0, n).astype(np.float32)
    tmp = np.zeros(n, np.float32)
    idx = rng.randint(0, n, 500).astype(np.uint32)
    val = rng.uniform(0, 2, 500).astype(np.float32)
    tmp2 = orc.splat_grid_samples(val, idx, tmp)
    ref = np.zeros(n, np.float32); np.maximum.at(ref, idx, val)
    assert np.array_equal(tmp2, ref)
    g2 = orc.ema_grid_samples_nerf(tmp2, 0.95, grid)
    exp = np.where(grid < 0, grid, np.maximum(grid * np.float32(0.95), tmp2))
    assert np.array_equal(g2, exp)
    bits = orc.grid_to_bitfield(g2, float(g2.mean()), 0.01)
    thr = min(np.float32(0.01), np.float32(g2.mean()))
    assert np.array_equal(np.unpackbits(bits, bitorder='little').astype(bool), g2 > thr)
    assert orc.count_bitfield(bits, faithful=False) == float((g2 > thr).sum())
    assert orc.count_bitfield(bits, faithful=True) == 8.0 * float((bits != 0).sum())
    grp = rng.randint(0, 50, 500)
    uni = orc.tensor_reduce_max(val, grp, 50)
    ref = np.zeros(50, np.float32); np.maximum.at(ref, grp, val)
    assert np.array_equal(uni, ref)


@pytest.mark.skipif(not have_ref(), reason='oracle/_ref not built (no /root/reference on this machine)')
def test_oracle_equals_reference_host_build_at_ngp_size(orc):
    """BASELINE configs[1] shape (4096 rays, 128^3, 1024 samples) straight against the reference functions."""
    ref = Sampler('ref')
    c = cases.march_case('ngp4096')
    a = orc.aabb_intersection(c['o'], c['d'], c['aabb'][None])
    b = ref.aabb_intersection(c['o'], c['d'], c['aabb'][None])
    for x, y in zip(a, b):
        assert np.array_equal(x, y)
    near, far = a[0], a[1]
    for kind in ('lego', 'core50'):
        vol = synth.occupancy_volume(kind)
        bits = synth.volume_to_bitfield(vol)
        for bit, fld in ((False, vol), (True, bits)):
            z1, m1 = orc.sparse_volume_sampling(c['o'], c['d'], near, far, c['n_pts'], c['dt'], c['aabb'], 128, fld,
                                                0.2, orc.rng(), bit=bit)
            z2, m2 = ref.sparse_volume_sampling(c['o'], c['d'], near, far, c['n_pts'], c['dt'], c['aabb'], 128, fld,
                                                0.2, ref.rng(), bit=bit)
            assert np.array_equal(m1, m2) and np.array_equal(z1, z2)
            assert m1.sum() > 50000
