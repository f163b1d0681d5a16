"""composite_train timing: real ray-length distribution vs uniform lengths with the same number of samples."""
import sys
import numpy as np, torch
sys.path.insert(0, '.')
from arcnerf_b200 import _lib, synth
from arcnerf_b200.ngp import NGPPipeline

n = 4096
pipe = NGPPipeline(max_rays=n, capacity=1 << 19, seed=1337)
pipe.set_occupancy(synth.occupancy_volume('lego'))
o, d = synth.train_ray_batch(n, seed=0)
o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
t = torch.rand((n, 3), device='cuda'); b = torch.rand((n, 3), device='cuda')
pipe.train_step(o, d, t, b)
torch.cuda.synchronize()
counts = pipe.counts[:n].cpu().numpy()
print('samples', counts.sum(), 'max', counts.max(), 'mean', counts.mean(), 'p50/p90/p99', np.percentile(counts, [50, 90, 99]), 'zero rays', (counts == 0).sum())
print('rays with >32: %.3f  >64: %.3f  >128: %.3f' % ((counts > 32).mean(), (counts > 64).mean(), (counts > 128).mean()))

def time_comp(offsets, N, pmax, reps=50):
    sig = torch.rand(N, device='cuda') * 20; rad = torch.rand((N, 3), device='cuda'); tt = torch.sort(torch.rand(N, device='cuda'))[0] + 2
    rgb = torch.empty((n, 3), device='cuda'); g = torch.empty((n, 3), device='cuda'); loss = torch.zeros(1, device='cuda')
    dg = torch.empty(N, device='cuda'); dr = torch.empty((N, 3), device='cuda')
    pm = torch.tensor([pmax], dtype=torch.int32, device='cuda')
    st = _lib.current_stream()
    def run():
        _lib.call('arcb200_composite_huber_train', offsets.data_ptr(), 0, N, pm.data_ptr(), sig.data_ptr(), rad.data_ptr(), tt.data_ptr(),
                  None, 0, 0, b.data_ptr(), 1, t.data_ptr(), 0.1, 3000.0, rgb.data_ptr(), g.data_ptr(), loss.data_ptr(), dg.data_ptr(),
                  dr.data_ptr(), None, 1024, n, st)
    for _ in range(5): run()
    torch.cuda._sleep(2000000)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): run()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3

off_real = torch.zeros(n + 1, dtype=torch.int32); off_real[1:] = torch.from_numpy(np.cumsum(counts)).int()
N = int(counts.sum())
print('real distribution      %.1f us' % time_comp(off_real.cuda(), N, int(counts.max())))
u = np.full(n, N // n); off_u = torch.zeros(n + 1, dtype=torch.int32); off_u[1:] = torch.from_numpy(np.cumsum(u)).int()
print('uniform %d per ray      %.1f us' % (N // n, time_comp(off_u.cuda(), int(u.sum()), int(u.max()))))
srt = np.sort(counts)[::-1].copy(); off_s = torch.zeros(n + 1, dtype=torch.int32); off_s[1:] = torch.from_numpy(np.cumsum(srt)).int()
print('real, sorted descending %.1f us' % time_comp(off_s.cuda(), N, int(counts.max())))
