"""Timing of the occupancy refresh (VolumeBound.optimize through the pipeline's density query)."""
import sys, time
import torch
sys.path.insert(0, '.')
from arcnerf_b200 import synth
from arcnerf_b200.ngp import NGPPipeline

pipe = NGPPipeline(max_rays=4096, capacity=1 << 19, seed=1337)
pipe.set_occupancy(synth.occupancy_volume('lego'))
for ep, tag in ((16, 'warm-up (2M voxels)'), (32, 'warm-up'), (256, 'steady (n/4 + n/4)'), (272, 'steady'), (288, 'steady')):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    pipe.update_occupancy(ep)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    print('epoch %4d %-22s %.3f ms   occupied %.3f' % (ep, tag, (t1 - t0) * 1e3, float(pipe.volume.float().mean())))
pts = torch.rand((1 << 20, 3), device='cuda') * 2 - 1
for _ in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    pipe.forward_pts(pts)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    print('forward_pts 1M points %.3f ms' % ((t1 - t0) * 1e3))
