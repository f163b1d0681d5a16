"""Refresh timing inside a training loop (events on the main stream)."""
import sys, time
import numpy as np, torch
sys.path.insert(0, '.')
from arcnerf_b200 import synth
from arcnerf_b200.ngp import NGPPipeline
n = 4096
pipe = NGPPipeline(max_rays=n, capacity=1 << 19, seed=1337)
pipe.set_occupancy(synth.occupancy_volume('lego'))
bat = []
for i in range(8):
    o, d = synth.train_ray_batch(n, seed=i)
    bat.append((torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda(), torch.rand((n, 3), device='cuda'), torch.rand((n, 3), device='cuda')))
pipe.update_occupancy(256, apply=False)
for i in range(20):
    pipe.train_step(*bat[i % 8])
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
t0 = time.perf_counter()
ev[0].record()
for i in range(16):
    pipe.train_step(*bat[i % 8], next_rays=bat[(i + 1) % 8][:2])
ev[1].record()
c0 = time.perf_counter()
pipe.update_occupancy(272, apply=False)
c1 = time.perf_counter()
ev[2].record()
for i in range(16):
    pipe.train_step(*bat[i % 8], next_rays=bat[(i + 1) % 8][:2])
ev[3].record()
torch.cuda.synchronize()
print('16 steps %.3f ms | refresh %.3f ms (cpu enqueue %.3f ms) | 16 steps %.3f ms' % (ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), (c1 - c0) * 1e3, ev[2].elapsed_time(ev[3])))
