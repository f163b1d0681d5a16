"""Timing of the dense-chain kernels at the vanilla-NeRF step size (1024 rays x 192 fine points = 196 608 points)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from arcnerf_b200 import _lib
from arcnerf_b200.plugin.nerf_mlp import FusedNeRFNets

def timeit(fn, reps=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return float(np.median(ts))

nets = FusedNeRFNets()
for n in (65536, 196608):
    x = torch.rand((n, 3), device='cuda') * 2 - 1
    d = torch.randn((n, 3), device='cuda')
    W, D, ex, ev, Wr = 256, 8, 63, 27, 128
    flops = 2.0 * n * (ex * W + (D - 1) * W * W + ex * W + W * (W + 1) + (ev + W) * Wr + Wr * 3)
    with torch.no_grad():
        t_inf = timeit(lambda: nets(x, d))
    sigma, rgb = nets(x, d)
    gs, gc = torch.randn_like(sigma), torch.randn_like(rgb)
    def fb():
        s, c = nets(x, d)
        torch.autograd.backward([s, c], [gs, gc])
    t_fb = timeit(fb)
    # separate kernels
    n_cap = n
    save = torch.empty(nets.fwd_cols * n_cap, dtype=torch.float16, device='cuda')
    sg = torch.empty(n, device='cuda'); rg = torch.empty((n, 3), device='cuda')
    nets._pack_if_changed()
    t_fwd_save = timeit(lambda: _lib.call('arcb200_nerf_mlp_fwd', *nets.arch, x.data_ptr(), d.data_ptr(), nets.blob_fwd.data_ptr(), nets.vec.data_ptr(),
                                          save.data_ptr(), n_cap, sg.data_ptr(), rg.data_ptr(), n, _lib.current_stream()))
    import os
    parts = {}
    for pp in ('1', '2'):
        os.environ['ARCB200_NERF_BWD_PARTS'] = pp
        parts[pp] = timeit(fb)
    os.environ.pop('ARCB200_NERF_BWD_PARTS')
    print('   fwd+save + dgrad chain only %.3f ms | fwd+save + wgrad only %.3f ms' % (parts['1'], parts['2']))
    print('n %d: forward (inference) %.3f ms = %.0f TFLOP/s | forward+save %.3f ms | forward+backward %.3f ms = %.0f TFLOP/s (3x forward flops)' % (
        n, t_inf, flops / t_inf / 1e9, t_fwd_save, t_fb, 3 * flops / t_fb / 1e9))

import ctypes as C
lib = _lib.load()
if hasattr(lib, 'arcb200_debug_chain_prof'):
    buf = (C.c_ulonglong * 32)()
    torch.cuda.synchronize(); lib.arcb200_debug_chain_prof(buf)
    with torch.no_grad():
        nets(x, d)
    torch.cuda.synchronize(); lib.arcb200_debug_chain_prof(buf)
    v = list(buf); pairs = max(v[6], 1)
    print('per pair of tiles (cycles): MMA thread waits weights %d, waits operand tiles %d, total %d | epilogue thread 0 waits accumulators %d, works %d | producer waits slot %d | pairs %d' % (
        v[0] / pairs, v[1] / pairs, v[2] / pairs, v[3] / pairs, v[4] / pairs, v[5] / pairs, pairs))
    print('epilogue work per layer (cycles/pair):', [int(v[8 + l] / pairs) for l in range(11)], 'load phase', int(v[24] / pairs), 'tmem wait total', int(v[25] / pairs))

if hasattr(lib, 'arcb200_debug_wg_prof'):
    b8 = (C.c_ulonglong * 8)()
    torch.cuda.synchronize(); lib.arcb200_debug_wg_prof(b8)
    fb(); torch.cuda.synchronize(); lib.arcb200_debug_wg_prof(b8)
    w = list(b8); it = max(w[5], 1)
    print('wgrad big jobs, cycles per 64-row iteration (thread 0): wait stage free %d, issue loads %d, wait loads %d, column sums %d, sync + MMA issue %d | iterations %d' % (
        w[0] / it, w[1] / it, w[2] / it, w[3] / it, w[4] / it, it))
