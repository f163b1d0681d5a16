"""Cameras and get_rays arguments of tests/golden/rays_golden.npz."""
import numpy as np

from arcnerf_b200 import synth


def cases():
    W, H = 40, 28
    K = np.array([[55.5, 0.3, 19.7], [0.0, 54.25, 13.6], [0.0, 0.0, 1.0]], np.float32)          # with skew and an off-centre principal point
    c2w = synth.look_at_pose((1.7, -2.1, 1.3)).astype(np.float32)
    idx = np.array([[0, 0], [39, 27], [5, 3], [17, 20], [39, 0]], np.int64)
    return {
        'full_wh': (W, H, K, c2w, dict(wh_order=True)),
        'full_hw_center': (W, H, K, c2w, dict(wh_order=False, center_pixel=True)),
        'full_unnormalised': (W, H, K, c2w, dict(normalize_rays_d=False)),
        'index': (W, H, K, c2w, dict(index=idx, center_pixel=True)),
    }
