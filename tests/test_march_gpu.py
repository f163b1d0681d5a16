"""GPU: ray/occupancy kernels through the C ABI (via the plugin wrappers) against the CPU oracle and the
golden vectors.  Bit-exact: sample distances, masks, counts, indices."""
import os

import numpy as np
import pytest
import torch

import cases
from arcnerf_b200 import _lib, synth
from arcnerf_b200.plugin import bitfield_func, multivol_func, volume_func
from oracle.sampler import Sampler

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'sampler_golden.npz'))


def T(a, dev, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    return t if dtype is None else t.to(dtype)


def range32(aabb, dev):
    """(2,3) [min;max] -> the reference's (3,2) aabb_range tensor"""
    return T(aabb.T.copy(), dev)


@pytest.fixture(scope='module')
def orc():
    return Sampler('oracle')


def test_aabb_table(cuda, orc):
    o, d = cases.aabb_table()
    aabb = np.stack([synth.aabb_of_side(2.0), np.array([[-0.5, -0.25, 0.0], [0.75, 1.0, 1.5]], np.float32)], 0)
    rng = torch.from_numpy(np.ascontiguousarray(aabb.transpose(0, 2, 1))).to(cuda)   # (N_v, 3, 2)
    near, far, pts, mask = volume_func.ray_aabb_intersection_cuda(T(o, cuda), T(d, cuda), rng)
    assert np.array_equal(mask.cpu().numpy(), GOLD['aabb_table_mask'])
    assert np.array_equal(near.cpu().numpy(), GOLD['aabb_table_near'])
    assert np.array_equal(far.cpu().numpy(), GOLD['aabb_table_far'])
    assert np.array_equal(pts.cpu().numpy(), GOLD['aabb_table_pts'], equal_nan=True)


@pytest.mark.parametrize('name', cases.MARCH_GOLDEN_CASES + ['ngp4096'])
def test_march_dense_vs_oracle(cuda, orc, name):
    c = cases.march_case(name)
    o, d = T(c['o'], cuda), T(c['d'], cuda)
    ar = range32(c['aabb'], cuda)
    near, far, _, mask_r = volume_func.ray_aabb_intersection_cuda(o, d, ar[None])
    near_o, far_o, _, mask_o = orc.aabb_intersection(c['o'], c['d'], c['aabb'][None])
    assert np.array_equal(near.cpu().numpy(), near_o) and np.array_equal(far.cpu().numpy(), far_o)
    assert np.array_equal(mask_r.cpu().numpy(), mask_o)
    bits = synth.volume_to_bitfield(c['vol'])
    volume_func.reset_rng()
    bitfield_func.reset_rng()
    bitfield_func._rng().advance()   # golden: the bit launch is the second of a session
    for bit, fld in ((0, c['vol']), (1, bits)):
        orng = orc.rng()
        if bit:
            orng.advance()
            z, m = bitfield_func.sparse_volume_sampling_bit(o, d, near, far, c['n_pts'], float(c['dt']), ar, c['n_grid'],
                                                            T(fld, cuda), c['near_distance'])
        else:
            z, m = volume_func.sparse_volume_sampling(o, d, near, far, c['n_pts'], float(c['dt']), ar, c['n_grid'],
                                                      T(fld, cuda), c['near_distance'])
        zo, mo = orc.sparse_volume_sampling(c['o'], c['d'], near_o, far_o, c['n_pts'], c['dt'], c['aabb'], c['n_grid'],
                                            fld, c['near_distance'], orng, bit=bool(bit))
        z, m = z.cpu().numpy(), m.cpu().numpy()
        assert m.dtype == np.bool_ and z.dtype == np.float32
        assert np.array_equal(m, mo)
        assert np.array_equal(z, zo)
        if name in cases.MARCH_GOLDEN_CASES:
            assert np.array_equal(m.sum(1).astype(np.int32), GOLD['march_%s_bit%d_counts' % (name, bit)])
            assert np.array_equal(z[m], GOLD['march_%s_bit%d_valid' % (name, bit)])
            assert np.array_equal(z[:, -1], GOLD['march_%s_bit%d_lastcol' % (name, bit)])
    # the module generators moved on by 2^32 per launch
    assert volume_func._rng().state == int(GOLD['rng_state_after_launch'][0])


def test_rng_advances_between_launches(cuda, orc):
    c = cases.march_case('small32')
    o, d, ar = T(c['o'], cuda), T(c['d'], cuda), range32(c['aabb'], cuda)
    near, far, _, _ = volume_func.ray_aabb_intersection_cuda(o, d, ar[None])
    vol = T(c['vol'], cuda)
    volume_func.reset_rng()
    z1, _ = volume_func.sparse_volume_sampling(o, d, near, far, c['n_pts'], float(c['dt']), ar, c['n_grid'], vol, 0.2)
    z2, _ = volume_func.sparse_volume_sampling(o, d, near, far, c['n_pts'], float(c['dt']), ar, c['n_grid'], vol, 0.2)
    assert not torch.equal(z1, z2)          # new jitter
    orng = orc.rng(); orng.advance()
    n_o, f_o, _, _ = orc.aabb_intersection(c['o'], c['d'], c['aabb'][None])
    zo, _ = orc.sparse_volume_sampling(c['o'], c['d'], n_o, f_o, c['n_pts'], c['dt'], c['aabb'], c['n_grid'], c['vol'],
                                       0.2, orng)
    assert np.array_equal(z2.cpu().numpy(), zo)


def test_check_pts_in_occ_voxel(cuda, orc):
    for name in cases.MARCH_GOLDEN_CASES:
        c = cases.march_case(name)
        pts = (c['o'][:, None] + c['d'][:, None] * np.linspace(0.5, 5.0, 24, dtype=np.float32)[None, :, None]).reshape(-1, 3)
        got = volume_func.check_pts_in_occ_voxel_cuda(T(pts, cuda), T(c['vol'], cuda), range32(c['aabb'], cuda), c['n_grid'])
        assert np.array_equal(got.cpu().numpy(), GOLD['occ_pts_%s' % name])


@pytest.mark.parametrize('inclusive', [False, True])
def test_multivol(cuda, inclusive):
    c = cases.multivol_case(inclusive)
    n = c['o'].shape[0]
    near = torch.zeros((n, 1), device=cuda)
    far = torch.full((n, 1), 12.0, device=cuda)
    multivol_func.reset_rng()
    z, m = multivol_func.sparse_sampling_in_multivol_bitfield(
        T(c['o'], cuda), T(c['d'], cuda), near, far, c['n_pts'], float(c['cone']), float(c['smin']), float(c['smax']),
        range32(c['min_aabb'], cuda), range32(c['aabb'], cuda), c['n_grid'], c['n_cascade'], T(c['bits'], cuda),
        c['near_distance'], inclusive)
    assert np.array_equal(m.cpu().numpy(), GOLD['multivol_incl%d_m' % inclusive])
    assert np.array_equal(z.cpu().numpy(), GOLD['multivol_incl%d_z' % inclusive])


def test_grid_maintenance(cuda, orc):
    rng = np.random.RandomState(31)
    grid = rng.uniform(-0.01, 0.05, size=(16 ** 3,)).astype(np.float32)
    bitfield_func.reset_rng()
    pos, idx = bitfield_func.generate_grid_samples(T(grid, cuda), 1000, 3, 16, 0.01)
    assert idx.dtype == torch.int32
    assert np.array_equal(idx.cpu().numpy().view(np.uint32), GOLD['gridsamples_idx'])
    assert np.array_equal(pos.cpu().numpy(), GOLD['gridsamples_pos'])
    gridm = rng.uniform(-0.01, 0.05, size=(16 ** 3 * 3,)).astype(np.float32)
    multivol_func.reset_rng()
    pos, idx = multivol_func.generate_grid_samples_multivol(T(gridm, cuda), 1000, range32(synth.aabb_of_side(2.0), cuda),
                                                            5, 4, 16, 0.01, False)
    assert np.array_equal(idx.cpu().numpy().view(np.uint32), GOLD['gridsamples_mv_idx'])
    assert np.array_equal(pos.cpu().numpy(), GOLD['gridsamples_mv_pos'])

    r = np.random.RandomState(4)
    n = 32 ** 3
    g = r.uniform(-0.5, 1.0, n).astype(np.float32)
    ix = r.randint(0, n, 5000).astype(np.int32)
    val = r.uniform(0, 2, 5000).astype(np.float32)
    tmp = bitfield_func.splat_grid_samples(T(val, cuda), T(ix, cuda), 5000, torch.zeros(n, device=cuda))
    tmp_o = orc.splat_grid_samples(val, ix.view(np.uint32), np.zeros(n, np.float32))
    assert np.array_equal(tmp.cpu().numpy(), tmp_o)
    g2 = bitfield_func.ema_grid_samples_nerf(tmp, T(g, cuda), n, 0.95)
    g2_o = orc.ema_grid_samples_nerf(tmp_o, 0.95, g)
    assert np.array_equal(g2.cpu().numpy(), g2_o)
    mean = float(g2_o.mean())
    bits = bitfield_func.update_bitfield(g2, mean, torch.empty(n // 8, dtype=torch.uint8, device=cuda), 0.01, 32)
    bits_o = orc.grid_to_bitfield(g2_o, mean, 0.01)
    assert np.array_equal(bits.cpu().numpy(), bits_o)
    bits_d = bitfield_func.update_bitfield(g2, g2.mean(), torch.empty(n // 8, dtype=torch.uint8, device=cuda), 0.01, 32)
    assert bits_d.sum().item() > 0
    assert bitfield_func.count_bitfield(bits, 32) == orc.count_bitfield(bits_o, faithful=True)
    assert bitfield_func.count_bitfield(bits, 32, faithful=False) == float(np.unpackbits(bits_o).sum())
    grp = r.randint(0, 50, 5000).astype(np.int64)
    uni = volume_func.tensor_reduce_max(T(val, cuda), T(grp, cuda), 50)
    assert np.array_equal(uni.cpu().numpy(), orc.tensor_reduce_max(val, grp, 50))


def test_march_packed_matches_dense(cuda, orc):
    for name in ['small32', 'full_cap', 'empty', 'ngp4096']:
        c = cases.march_case(name)
        n = c['o'].shape[0]
        o, d, ar = T(c['o'], cuda), T(c['d'], cuda), range32(c['aabb'], cuda)
        aabb23 = T(c['aabb'], cuda)
        near, far, _, _ = volume_func.ray_aabb_intersection_cuda(o, d, ar[None])
        vol = T(c['vol'], cuda)
        volume_func.reset_rng()
        z, m = volume_func.sparse_volume_sampling(o, d, near, far, c['n_pts'], float(c['dt']), ar, c['n_grid'], vol,
                                                  c['near_distance'])
        cap = int(m.sum().item()) + 7
        nbytes = _lib.load().arcb200_march_packed_scratch_bytes(n, c['n_pts'])
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=cuda)
        counts = torch.empty(n, dtype=torch.int32, device=cuda)
        offsets = torch.empty(n + 1, dtype=torch.int32, device=cuda)
        meta = torch.empty(2, dtype=torch.int32, device=cuda)
        t = torch.full((cap,), -1.0, device=cuda)
        xyz = torch.zeros((cap, 3), device=cuda)
        rid = torch.full((cap,), -1, dtype=torch.int32, device=cuda)
        unit = torch.zeros((cap, 3), device=cuda)
        r = _lib.Rng()
        _lib.call('arcb200_march_packed', 0, o.data_ptr(), d.data_ptr(), near.data_ptr(), far.data_ptr(), c['n_pts'],
                  float(c['dt']), aabb23.data_ptr(), c['n_grid'], vol.data_ptr(), float(c['near_distance']), n,
                  r.state, r.inc, scratch.data_ptr(), counts.data_ptr(), offsets.data_ptr(), meta.data_ptr(), cap,
                  t.data_ptr(), xyz.data_ptr(), rid.data_ptr(), unit.data_ptr(), _lib.current_stream())
        cnt = m.sum(1).to(torch.int32)
        assert torch.equal(counts, cnt)
        total = int(cnt.sum().item())
        assert meta.tolist() == [total, int(cnt.max().item())]
        exp_off = torch.zeros(n + 1, dtype=torch.int64, device=cuda)
        exp_off[1:] = torch.cumsum(cnt.long(), 0)
        assert torch.equal(offsets.long(), exp_off)
        assert torch.equal(t[:total], z[m])                       # ray-major order == boolean-mask order
        assert torch.all(t[total:] == -1.0)
        rid_exp = torch.repeat_interleave(torch.arange(n, device=cuda, dtype=torch.int32), cnt.long())
        assert torch.equal(rid[:total], rid_exp)
        if total:
            # positions: o + d*t with separately rounded multiply and add (the oracle's ray_at)
            oo, dd, tt = c['o'][rid_exp.cpu().numpy()], c['d'][rid_exp.cpu().numpy()], t[:total].cpu().numpy()
            exp = (oo + (dd * tt[:, None]).astype(np.float32)).astype(np.float32)
            assert np.array_equal(xyz[:total].cpu().numpy(), exp)
            lo, hi = c['aabb'][0], c['aabb'][1]
            assert np.array_equal(unit[:total].cpu().numpy(), ((exp - lo) / (hi - lo)).astype(np.float32))
