"""Device ray generation (csrc/raygen.cu, plugin/ray_helper.get_rays) against the REFERENCE's own get_rays run on CPU
(tests/golden/rays_golden.npz).  Same expressions in the same order in fp32: directions agree to 2e-6 (division and sqrt
rounding of the two libraries), origins exactly."""
import os

import numpy as np
import pytest
import torch

import ray_cases as rcs

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'rays_golden.npz'))


@pytest.mark.parametrize('name', list(rcs.cases()))
def test_get_rays_matches_reference(name):
    from arcnerf_b200.plugin import ray_helper as rh
    W, H, K, c2w, kw = rcs.cases()[name]
    o, d, idx, r = rh.get_rays(W, H, torch.from_numpy(K).cuda(), torch.from_numpy(c2w).cuda(), **kw)
    assert np.array_equal(o.cpu().numpy(), G[name + '/rays_o'])
    assert np.abs(d.cpu().numpy() - G[name + '/rays_d']).max() < 2e-6 * max(1.0, np.abs(G[name + '/rays_d']).max())
    if name + '/rays_r' in G.files:
        assert r.shape == (W * H, 1) and np.abs(r.cpu().numpy() - G[name + '/rays_r']).max() < 1e-6
    else:
        assert r is None
    if name + '/index' in G.files:
        assert idx == G[name + '/index'].tolist()


def test_batch_from_many_cameras_in_one_launch():
    from arcnerf_b200 import synth
    from arcnerf_b200.plugin import ray_helper as rh
    W, H = 64, 48
    K = np.array([[70.0, 0.0, 32.0], [0.0, 70.0, 24.0], [0.0, 0.0, 1.0]], np.float32)
    poses = synth.sphere_poses(5, seed=2).astype(np.float32)
    cams = rh.pack_cameras(np.repeat(K[None], 5, 0), poses, 'cuda')
    rs = np.random.RandomState(0)
    cam_id = torch.from_numpy(rs.randint(0, 5, 500).astype(np.int32)).cuda()
    pix = torch.from_numpy(rs.randint(0, W * H, 500).astype(np.int64)).cuda()
    o, d = rh.get_train_batch_rays(cams, cam_id, pix, W, H, center_pixel=True)
    for c in range(5):
        sel = (cam_id == c).nonzero()[:, 0]
        idx2 = torch.stack([pix[sel] // H, pix[sel] % H], 1)
        o1, d1, _, _ = rh.get_rays(W, H, torch.from_numpy(K).cuda(), torch.from_numpy(poses[c]).cuda(), index=idx2, center_pixel=True)
        assert torch.equal(o[sel], o1) and torch.equal(d[sel], d1)


def test_bad_arguments():
    from arcnerf_b200 import _lib
    with pytest.raises(RuntimeError):
        _lib.call('arcb200_gen_rays', 1, 1, None, None, 4, 4, 1, 0, 1, 1, 1, None, 7, None)      # n != W*H without a pixel list
