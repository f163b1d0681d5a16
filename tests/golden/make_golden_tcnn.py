"""Writes tests/golden/tcnn_golden.npz: outputs of the REFERENCE's own tiny-cuda-nn (ArcNerf's vendored submodule,
/root/reference/include/tiny-cuda-nn @ 8e6e242f, built for sm_100 into oracle/_ref/libtcnnref.so by
`make -C oracle tcnn`) for the seeded inputs of tests/tcnn_cases.py.  Needs a GPU: run on the B200 box,

    gpurun -- python tests/golden/make_golden_tcnn.py gpurun_out/tcnn_golden.npz

then move the file to tests/golden/.  The call protocol is the reference's python layer (tinycudann/modules.py:58-159):
parameters cast to fp16, batches of a multiple of 128 rows, gradients w.r.t. the output multiplied by loss_scale = 128
before the native backward and the results divided by it afterwards.
"""
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import tcnn_cases as tc          # noqa: E402

LOSS_SCALE = 128.0


class Tcnn:
    def __init__(self):
        self.lib = C.CDLL(os.path.join(ROOT, 'oracle', '_ref', 'libtcnnref.so'))
        L = self.lib
        for name in ('tcnnref_create_encoding', 'tcnnref_create_network', 'tcnnref_forward'):
            getattr(L, name).restype = C.c_void_p
        L.tcnnref_last_error.restype = C.c_char_p
        L.tcnnref_n_params.restype = C.c_longlong

    def err(self):
        return self.lib.tcnnref_last_error().decode()

    def encoding(self, n_in, cfg, fp16=True):
        m = self.lib.tcnnref_create_encoding(n_in, json.dumps(cfg).encode(), int(fp16))
        assert m, self.err()
        return Module(self, m)

    def network(self, n_in, n_out, cfg):
        m = self.lib.tcnnref_create_network(n_in, n_out, json.dumps(cfg).encode())
        assert m, self.err()
        return Module(self, m)


class Module:
    def __init__(self, t, handle):
        self.t, self.lib, self.h = t, t.lib, C.c_void_p(handle)
        self.n_params = int(self.lib.tcnnref_n_params(self.h))
        self.n_in = self.lib.tcnnref_n_input_dims(self.h)
        self.n_out = self.lib.tcnnref_n_output_dims(self.h)
        self.p16 = bool(self.lib.tcnnref_param_is_fp16(self.h))
        self.o16 = bool(self.lib.tcnnref_output_is_fp16(self.h))
        self.pdt = torch.float16 if self.p16 else torch.float32
        self.odt = torch.float16 if self.o16 else torch.float32

    def initial_params(self, seed=1337):
        p = torch.zeros(self.n_params, dtype=torch.float32, device='cuda')
        assert self.lib.tcnnref_initial_params(self.h, C.c_ulonglong(seed), C.c_void_p(p.data_ptr())) == 0, self.t.err()
        torch.cuda.synchronize()
        return p

    def forward(self, x, params, input_grad=False):
        n = x.shape[0]
        assert n % 128 == 0 and x.dtype == torch.float32 and params.dtype == self.pdt
        out = torch.empty((n, self.n_out), dtype=self.odt, device='cuda')
        ctx = self.lib.tcnnref_forward(self.h, None, n, C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()),
                                       C.c_void_p(params.data_ptr()), int(input_grad))
        assert ctx, self.t.err()
        return C.c_void_p(ctx), out

    def inference(self, x, params):
        n = x.shape[0]
        out = torch.empty((n, self.n_out), dtype=self.odt, device='cuda')
        assert self.lib.tcnnref_inference(self.h, None, n, C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()),
                                          C.c_void_p(params.data_ptr())) == 0, self.t.err()
        return out

    def backward(self, ctx, x, params, out, dy, want_dx=True, want_dp=True):
        """modules.py:86-97: dy * loss_scale in, results / loss_scale out"""
        n = x.shape[0]
        sdy = (dy * LOSS_SCALE).to(self.odt).contiguous()
        dx = torch.empty((n, self.n_in), dtype=torch.float32, device='cuda') if want_dx else None
        dp = torch.empty(self.n_params, dtype=self.pdt, device='cuda') if want_dp else None
        p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
        assert self.lib.tcnnref_backward(self.h, None, ctx, n, p(dx), p(sdy), p(dp), p(x), p(out), p(params)) == 0, self.t.err()
        torch.cuda.synchronize()
        return (dx / LOSS_SCALE if dx is not None else None), (dp / LOSS_SCALE if dp is not None else None)

    def backward_backward_input(self, ctx, x, params, v, dy):
        """modules.py:99-127"""
        n = x.shape[0]
        sdy = (dy * LOSS_SCALE).to(self.odt).contiguous()
        d_dy = torch.zeros((n, self.n_out), dtype=self.odt, device='cuda')
        dp = torch.zeros(self.n_params, dtype=self.pdt, device='cuda')
        dx = torch.zeros((n, self.n_in), dtype=torch.float32, device='cuda')
        p = lambda t: C.c_void_p(t.data_ptr())
        assert self.lib.tcnnref_backward_backward_input(self.h, None, ctx, n, p(v), p(x), p(sdy), p(dp), p(d_dy), p(dx),
                                                        p(params)) == 0, self.t.err()
        torch.cuda.synchronize()
        return d_dy, dp / LOSS_SCALE, dx / LOSS_SCALE

    def free(self, ctx):
        self.lib.tcnnref_free_context(ctx)


def sparse(t):
    """flat gradient -> (indices int32, values) of its non-zero entries"""
    t = t.detach().float().cpu().numpy()
    idx = np.flatnonzero(t).astype(np.int32)
    return idx, t[idx].astype(np.float16 if np.all(np.abs(t[idx]) < 6e4) else np.float32)


def main(out_path):
    assert torch.cuda.is_available()
    T = Tcnn()
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    G = {}
    G['meta_device'] = np.array(torch.cuda.get_device_name(0))
    G['meta_batch_granularity'] = np.array(T.lib.tcnnref_batch_granularity())

    # ---------------------------------------------------------------- hash grid (hashgrid_encoder.py:102-113)
    g = tc.GRID
    pls = tc.per_level_scale()
    enc = T.encoding(3, {'otype': 'HashGrid', 'n_levels': g['n_levels'], 'n_features_per_level': g['n_features'],
                         'log2_hashmap_size': g['log2_hashmap_size'], 'base_resolution': g['base_res'],
                         'per_level_scale': pls})
    G['grid_n_params'] = np.array(enc.n_params)
    # the per-level scales exactly as tcnn's kernels compute them on the device (exp2f there is not the host's)
    sc = torch.zeros(g['n_levels'], dtype=torch.float32, device='cuda')
    T.lib.tcnnref_device_grid_scales(g['n_levels'], C.c_float(float(np.log2(np.float32(pls)))), g['base_res'], C.c_void_p(sc.data_ptr()))
    torch.cuda.synchronize()
    G['grid_device_scales'] = sc.cpu().numpy()
    G['grid_per_level_scale'] = np.array(pls)
    assert enc.p16 and enc.o16 and enc.n_out == 32
    init = enc.initial_params(1337)
    G['grid_init_head'] = init[:4096].cpu().numpy()
    G['grid_init_stats'] = np.array([float(init.min()), float(init.max()), float(init.mean()), float(init.std())])
    table = cu(tc.grid_table(enc.n_params)).half()
    x = cu(tc.grid_points())
    dy = cu(tc.grid_dy())
    v = cu(tc.grid_v())
    ctx, y = enc.forward(x, table, input_grad=True)
    torch.cuda.synchronize()
    G['grid_fwd'] = y.cpu().numpy()
    G['grid_inference'] = enc.inference(x, table).cpu().numpy()
    dx, dp = enc.backward(ctx, x, table, y, dy)
    G['grid_dLdx'] = dx.cpu().numpy()
    G['grid_dLdp_idx'], G['grid_dLdp_val'] = sparse(dp)
    d_dy, dp2, dx2 = enc.backward_backward_input(ctx, x, table, v, dy)
    G['grid_bb_dLddLdy'] = d_dy.cpu().numpy()
    G['grid_bb_dLdp_idx'], G['grid_bb_dLdp_val'] = sparse(dp2)
    G['grid_bb_dLdx'] = dx2.cpu().numpy()
    enc.free(ctx)

    # ---------------------------------------------------------------- spherical harmonics (sh_encoder.py:59-65,116-118)
    sh = T.encoding(3, {'otype': 'SphericalHarmonics', 'degree': 4})
    d = cu(tc.sh_dirs())
    dn = (d + 1) / 2.0
    dummy = torch.zeros(max(sh.n_params, 1), dtype=sh.pdt, device='cuda')
    G['sh_fwd'] = sh.inference(dn.contiguous(), dummy).cpu().numpy()

    # ---------------------------------------------------------------- FullyFusedMLP (tcnn_fusedmlp_module.py:66-77,162-173)
    for k, (name, in_dim, width, n_hidden, out_dim, act, out_act) in enumerate(tc.MLPS):
        net = T.network(in_dim, out_dim, {'otype': 'FullyFusedMLP', 'activation': act, 'output_activation': out_act,
                                          'n_neurons': width, 'n_hidden_layers': n_hidden})
        assert net.n_params == tc.mlp_n_params(in_dim, width, n_hidden, out_dim), (net.n_params, name)
        G['mlp_%s_n_out' % name] = np.array(net.n_out)
        G['mlp_%s_init' % name] = net.initial_params(1337).cpu().numpy()
        params = cu(tc.mlp_params(in_dim, width, n_hidden, out_dim, seed=200 + k)).half()
        xin = cu(tc.mlp_inputs(in_dim, seed=300 + k))
        ctx, y = net.forward(xin, params, input_grad=True)
        torch.cuda.synchronize()
        G['mlp_%s_fwd' % name] = y.cpu().numpy()
        G['mlp_%s_inference' % name] = net.inference(xin, params).cpu().numpy()
        net.free(ctx)
        # backward on gate-stable rows (see tcnn_cases.mlp_inputs_gate_stable)
        xs = cu(tc.mlp_inputs_gate_stable(in_dim, width, n_hidden, out_dim, params.float().cpu().numpy(), seed=500 + k))
        ctx, ys = net.forward(xs, params, input_grad=True)
        torch.cuda.synchronize()
        G['mlp_%s_stable_fwd' % name] = ys.cpu().numpy()
        gy = cu(tc.mlp_dy(net.n_out, out_dim, seed=400 + k))
        dx, dp = net.backward(ctx, xs, params, ys, gy)
        G['mlp_%s_dLdx' % name] = dx.cpu().numpy()
        G['mlp_%s_dLdp' % name] = dp.float().cpu().numpy()
        net.free(ctx)

    os.makedirs(os.path.dirname(os.path.abspath(out_path)), exist_ok=True)
    np.savez_compressed(out_path, **G)
    print('wrote', out_path, {k: getattr(a, 'shape', None) for k, a in G.items()})


if __name__ == '__main__':
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, 'gpurun_out', 'tcnn_golden.npz'))
