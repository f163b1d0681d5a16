"""GPU-box script (test infrastructure / profiling, not the product path): runs the REFERENCE's own kernels, built for sm_100
by `make -C oracle ref_dev tcnn`, next to this library's on identical inputs and writes gpurun_out/r02_reference_kernels.json:
  * arcnerf/ops device build (nvcc default: FMA contraction on) vs the pinned canonical samples (the reference's host build
    semantics = oracle/sampler_oracle.c, no contraction): per-case mismatch rates of sample counts and distances;
  * timings (CUDA events, best of 20 after 5 warm-ups) of the reference's A1 / A2 / B1 / C1 kernels and of tcnn's hash grid
    and FullyFusedMLP forward / backward, beside the same op of libarcb200.
"""
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
sys.path.insert(0, os.path.join(ROOT, 'tests', 'golden'))

import cases                                    # noqa: E402
import tcnn_cases as tc                          # noqa: E402
from arcnerf_b200 import _lib, synth            # noqa: E402
from arcnerf_b200 import functional as F         # noqa: E402
from oracle import sampler as osamp              # noqa: E402

cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
P = lambda t: C.c_void_p(t.data_ptr())


def timeit(fn, reps=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


def sampler_part(out):
    ref = {m: C.CDLL(os.path.join(ROOT, 'oracle', '_ref', 'libarcrefdev_%s.so' % m)) for m in ('volume', 'bitfield', 'multivol')}
    orc = osamp.Sampler('oracle')
    res = {}
    for name in cases.MARCH_GOLDEN_CASES + ['ngp4096']:
        c = cases.march_case(name)
        o, d, aabb = cu(c['o']), cu(c['d']), cu(c['aabb'])
        n = o.shape[0]
        near = torch.zeros((n, 1), device='cuda'); far = torch.zeros((n, 1), device='cuda')
        pts = torch.zeros((n, 1, 2, 3), device='cuda'); hit = torch.zeros((n, 1), dtype=torch.uint8, device='cuda')
        ref['volume'].arcrefdev_aabb_intersection(n, 1, P(o), P(d), P(aabb), P(near), P(far), P(pts), P(hit))
        torch.cuda.synchronize()
        on, of, _, om = orc.aabb_intersection(c['o'], c['d'], c['aabb'][None])
        r = dict(rays=n, aabb_near_mismatch=float((near.cpu().numpy().view(np.uint32) != on.view(np.uint32)).mean()),
                 aabb_far_mismatch=float((far.cpu().numpy().view(np.uint32) != of.view(np.uint32)).mean()),
                 aabb_mask_mismatch=float((hit.cpu().numpy().astype(bool) != om).mean()))
        # canonical samples from the canonical near/far (what the pinned pipeline feeds the march)
        nr, fr = cu(on), cu(of)
        vol = cu(c['vol'].astype(np.uint8))
        bits = cu(synth.volume_to_bitfield(c['vol']))
        for mod, fn, fld, bit in (('volume', 'arcrefdev_sparse_volume_sampling', vol, False),
                                  ('bitfield', 'arcrefdev_sparse_volume_sampling_bit', bits, True)):
            z = torch.zeros((n, c['n_pts']), device='cuda'); m = torch.zeros((n, c['n_pts']), dtype=torch.uint8, device='cuda')
            ref[mod].arcrefdev_reset_rng()
            getattr(ref[mod], fn)(n, P(o), P(d), P(nr), P(fr), c['n_pts'], C.c_float(c['dt']), P(aabb), c['n_grid'], P(fld),
                                  C.c_float(c['near_distance']), P(z), P(m))
            torch.cuda.synchronize()
            oz, omk = orc.sparse_volume_sampling(c['o'], c['d'], on, of, c['n_pts'], c['dt'], c['aabb'], c['n_grid'],
                                                 synth.volume_to_bitfield(c['vol']) if bit else c['vol'], c['near_distance'],
                                                 orc.rng(), bit=bit)
            zc, mc = z.cpu().numpy(), m.cpu().numpy().astype(bool)
            cnt_ref, cnt_orc = mc.sum(1), omk.sum(1)
            both = mc & omk
            r[mod] = dict(rays_with_different_count=float((cnt_ref != cnt_orc).mean()),
                          samples_ref=int(cnt_ref.sum()), samples_canonical=int(cnt_orc.sum()),
                          zvals_bit_mismatch_among_common=float((zc.view(np.uint32)[both] != oz.view(np.uint32)[both]).mean()) if both.any() else 0.0,
                          max_abs_zval_diff=float(np.abs(zc[both] - oz[both]).max()) if both.any() else 0.0)
            if name == 'ngp4096':
                def run():
                    getattr(ref[mod], fn)(n, P(o), P(d), P(nr), P(fr), c['n_pts'], C.c_float(c['dt']), P(aabb), c['n_grid'], P(fld),
                                          C.c_float(c['near_distance']), P(z), P(m))
                r[mod]['reference_kernel_ms_incl_devsync'] = timeit(run)
        if name == 'ngp4096':
            r['aabb_reference_kernel_ms_incl_devsync'] = timeit(lambda: ref['volume'].arcrefdev_aabb_intersection(
                n, 1, P(o), P(d), P(aabb), P(near), P(far), P(pts), P(hit)))
            # this library: slab test + packed march (4 launches) on the same rays
            lib = _lib.load()
            counts = torch.empty(n, dtype=torch.int32, device='cuda'); offs = torch.empty(n + 1, dtype=torch.int32, device='cuda')
            meta = torch.zeros(2, dtype=torch.int32, device='cuda')
            scratch = torch.empty(lib.arcb200_march_packed_scratch_bytes(n, c['n_pts']), dtype=torch.uint8, device='cuda')
            cap = 1 << 19
            t = torch.empty(cap, device='cuda'); unit = torch.empty((cap, 3), device='cuda'); rid = torch.empty(cap, dtype=torch.int32, device='cuda')
            rng = _lib.Rng()
            volb = cu(c['vol'])
            st = _lib.current_stream()

            def ours():
                _lib.call('arcb200_aabb_intersection', o.data_ptr(), d.data_ptr(), aabb.data_ptr(), near.data_ptr(), far.data_ptr(),
                          None, hit.data_ptr(), n, 1, st)
                _lib.call('arcb200_march_packed', 0, o.data_ptr(), d.data_ptr(), near.data_ptr(), far.data_ptr(), c['n_pts'],
                          float(c['dt']), aabb.data_ptr(), c['n_grid'], volb.data_ptr(), c['near_distance'], n, rng.state, rng.inc,
                          scratch.data_ptr(), counts.data_ptr(), offs.data_ptr(), meta.data_ptr(), cap, t.data_ptr(), None,
                          rid.data_ptr(), unit.data_ptr(), st)
            r['arcb200_aabb_plus_packed_march_ms'] = timeit(ours)
        res[name] = r
    # cascaded march
    for inclusive in (False, True):
        c = cases.multivol_case(inclusive)
        n = c['o'].shape[0]
        o, d = cu(c['o']), cu(c['d'])
        near = np.full((n, 1), 0.0, np.float32); far = np.full((n, 1), 20.0, np.float32)
        z = torch.zeros((n, c['n_pts']), device='cuda'); m = torch.zeros((n, c['n_pts']), dtype=torch.uint8, device='cuda')
        bits = cu(c['bits'])
        near_d, far_d, mab_d, ab_d = cu(near), cu(far), cu(c['min_aabb']), cu(c['aabb'])     # keep the device copies alive
        ref['multivol'].arcrefdev_reset_rng()
        ref['multivol'].arcrefdev_sparse_sampling_in_multivol_bitfield(
            n, P(o), P(d), P(near_d), P(far_d), c['n_pts'], C.c_float(c['cone']), C.c_float(c['smin']), C.c_float(c['smax']),
            P(mab_d), P(ab_d), c['n_grid'], c['n_cascade'], P(bits), C.c_int64(bits.numel()),
            C.c_float(c['near_distance']), int(inclusive), P(z), P(m))
        torch.cuda.synchronize()
        oz, omk = orc.sparse_sampling_in_multivol_bitfield(c['o'], c['d'], near, far, c['n_pts'], c['cone'], c['smin'], c['smax'],
                                                           c['min_aabb'], c['aabb'], c['n_grid'], c['n_cascade'], c['bits'],
                                                           c['near_distance'], inclusive, orc.rng())
        zc, mc = z.cpu().numpy(), m.cpu().numpy().astype(bool)
        both = mc & omk
        res['multivol_inclusive_%d' % int(inclusive)] = dict(
            rays=n, rays_with_different_count=float((mc.sum(1) != omk.sum(1)).mean()),
            zvals_bit_mismatch_among_common=float((zc.view(np.uint32)[both] != oz.view(np.uint32)[both]).mean()) if both.any() else 0.0)
    out['arcnerf_ops_device_build_vs_canonical'] = res


def tcnn_part(out):
    from make_golden_tcnn import Tcnn
    T = Tcnn()
    g = tc.GRID
    n = 147712                                       # the bench's ~147.6k samples per step, rounded up to 128
    enc = T.encoding(3, {'otype': 'HashGrid', 'n_levels': g['n_levels'], 'n_features_per_level': g['n_features'],
                         'log2_hashmap_size': g['log2_hashmap_size'], 'base_resolution': g['base_res'],
                         'per_level_scale': tc.per_level_scale()})
    table32 = cu(tc.grid_table(enc.n_params))
    table = table32.half()
    x = cu(np.random.RandomState(1).uniform(0, 1, (n, 3)).astype(np.float32))
    dy = cu(np.random.RandomState(2).normal(0, 1, (n, 32)).astype(np.float16))
    res = {'rows': n}
    res['tcnn_param_cast_fp32_to_fp16_ms (every forward, modules.py:156)'] = timeit(lambda: table32.half())
    res['tcnn_grid_inference_ms'] = timeit(lambda: enc.inference(x, table))
    ctx, y = enc.forward(x, table, input_grad=False)
    sdy = (dy * 128.0).contiguous()
    dp = torch.empty(enc.n_params, dtype=torch.float16, device='cuda')

    def grid_bwd():
        assert enc.lib.tcnnref_backward(enc.h, None, ctx, n, None, P(sdy), P(dp), P(x), P(y), P(table)) == 0
    res['tcnn_grid_backward_params_ms'] = timeit(grid_bwd)
    geom = F.GridGeometry(g['n_levels'], g['n_features'], g['log2_hashmap_size'], g['base_res'], tc.per_level_scale())
    enc_out = torch.empty((n, 32), dtype=torch.float16, device='cuda')
    res['arcb200_grid_fwd_ms'] = timeit(lambda: F.grid_forward(x, table, geom, out=enc_out))
    gt = torch.zeros(geom.n_params, device='cuda')
    dyf = dy.float()
    res['arcb200_grid_bwd_ms'] = timeit(lambda: F.grid_backward(x, dyf, geom, grad_table=gt))
    for k, (name, in_dim, width, n_hidden, out_dim, act, out_act) in enumerate(tc.MLPS):
        net = T.network(in_dim, out_dim, {'otype': 'FullyFusedMLP', 'activation': act, 'output_activation': out_act,
                                          'n_neurons': width, 'n_hidden_layers': n_hidden})
        params32 = cu(tc.mlp_params(in_dim, width, n_hidden, out_dim, seed=200 + k))
        params = params32.half()
        xin = cu(np.random.RandomState(5 + k).uniform(-1, 1, (n, in_dim)).astype(np.float32))
        res['tcnn_mlp_%s_inference_ms' % name] = timeit(lambda: net.inference(xin, params))
        c2, y2 = net.forward(xin, params, input_grad=True)
        gy = cu(tc.mlp_dy(net.n_out, out_dim, n=n, seed=9)) * 128.0
        dxi = torch.empty((n, in_dim), device='cuda'); dpp = torch.empty(net.n_params, dtype=torch.float16, device='cuda')

        def mlp_bwd():
            assert net.lib.tcnnref_backward(net.h, None, c2, n, P(dxi), P(gy), P(dpp), P(xin), P(y2), P(params)) == 0
        res['tcnn_mlp_%s_backward_ms' % name] = timeit(mlp_bwd)
        spec = F.MlpSpec(in_dim, width, n_hidden, out_dim, act.lower(), out_act.lower())
        blob = F.mlp_pack(params32, spec)
        xh = xin.half()
        yo = torch.empty((n, out_dim), device='cuda')
        res['arcb200_mlp_%s_fwd_ms' % name] = timeit(lambda: F.mlp_forward(xh, blob, spec, out=yo))
        dW = torch.zeros(spec.n_params, device='cuda'); dxo = torch.empty((n, in_dim), device='cuda')
        gyo = gy[:, :out_dim].float().contiguous()
        res['arcb200_mlp_%s_bwd_ms' % name] = timeit(lambda: F.mlp_backward(xh, blob, spec, gyo, dW=dW, dx=dxo))
    out['tcnn_vs_arcb200_kernels'] = res


if __name__ == '__main__':
    out = {'device': torch.cuda.get_device_name(0)}
    for part in (sampler_part, tcnn_part):
        try:
            part(out)
        except Exception as e:                     # keep what was measured
            import traceback
            out[part.__name__ + '_error'] = traceback.format_exc()
    os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
    with open(os.path.join(ROOT, 'gpurun_out', 'r02_reference_kernels.json'), 'w') as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))
