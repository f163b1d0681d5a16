"""Phase clocks of the fused field forward (debug build: ARCB200_NVCC_EXTRA=-DARCB200_PROF python -m arcnerf_b200.build --force)."""
import ctypes as C, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from arcnerf_b200 import _lib, synth
from arcnerf_b200.ngp import NGPPipeline
lib = _lib.load()
pipe = NGPPipeline(max_rays=32768, capacity=3 << 19)
pipe.set_occupancy(synth.occupancy_volume('lego'))
names = ['tiles', 'wait_tile', '-', 'geo0', 'geo1', 'rad0', 'rad1', 'rad2', 'layers_total', 'g_wait_free', 'g_gather', 'g_total']
for label, n_rays, train in (('train 4096 rays', 4096, True), ('render 32768 rays', 32768, False)):
    o, d = synth.train_ray_batch(n_rays, seed=1)
    o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    pipe.training = train
    for it in range(3):
        pipe.sample(o, d)
        torch.cuda.synchronize()
        buf = (C.c_ulonglong * 16)()
        lib.arcb200_debug_field_prof(buf)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); pipe.field_forward(d); b.record(); torch.cuda.synchronize()
        lib.arcb200_debug_field_prof(buf)
    v = np.array(list(buf), np.float64)
    print(label, 'samples', int(pipe.meta[0]), 'kernel ms %.4f' % a.elapsed_time(b), 'tiles', int(v[0]))
    for k in range(1, 12):
        print('   %-10s %8.0f cycles/tile' % (names[k], v[k] / max(v[0], 1)))
