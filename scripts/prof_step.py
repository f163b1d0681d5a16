"""A few eager training steps of the NGP pipeline at the bench's size, for ncu captures (kernel filter on the ncu side)."""
import sys
import numpy as np
import torch
sys.path.insert(0, '.')
from arcnerf_b200 import synth
from arcnerf_b200.ngp import NGPPipeline

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 4
n = 4096
pipe = NGPPipeline(max_rays=n, capacity=1 << 19, device='cuda:0', seed=1337)
pipe.set_occupancy(synth.occupancy_volume('lego'))
for i in range(steps):
    o, d = synth.train_ray_batch(n, seed=i)
    rng = np.random.RandomState(1000 + i)
    t, b = (torch.from_numpy(rng.uniform(0, 1, (n, 3)).astype(np.float32)).cuda() for _ in range(2))
    loss = pipe.train_step(torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda(), t, b)
torch.cuda.synchronize()
print('loss', float(loss))
