"""Tensor-level host functions over the C ABI (include/arcb200.h): each takes/returns torch CUDA tensors, allocates
the outputs with torch (plumbing only) and enqueues exactly the library kernels on torch's current stream.
The nn.Module mirrors in arcnerf_b200/plugin and the fused NGP pipeline are built from these.
No function here has a CPU or PyTorch fallback: a missing library or a failed call raises RuntimeError.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib

ACT_CODES = {None: 0, 'none': 0, 'None': 0, 'relu': 1, 'ReLU': 1, 'sigmoid': 2, 'Sigmoid': 2, 'softplus': 3,
             'Softplus': 3, 'exp': 4, 'Exponential': 4}


def _cuda(t, name):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError('%s must be a CUDA tensor' % name)
    if not t.is_contiguous():
        raise RuntimeError('%s must be contiguous' % name)
    return t


def _stream():
    return _lib.current_stream()


# ---------------------------------------------------------------------------------------------- hash grid
class GridGeometry:
    """Per-level geometry of a tcnn-style HashGrid, computed once on the host by the library
    (arcb200_grid_geometry; tcnn encodings/grid.h:195-204,982-1012)."""

    def __init__(self, n_levels=16, n_features=2, log2_hashmap_size=19, base_resolution=16, per_level_scale=2.0):
        self.n_levels, self.n_features = int(n_levels), int(n_features)
        self.log2_hashmap_size, self.base_resolution = int(log2_hashmap_size), int(base_resolution)
        self.per_level_scale = float(per_level_scale)
        self.offsets = np.zeros(self.n_levels + 1, np.uint32)
        self.scales = np.zeros(self.n_levels, np.float32)
        self.resolutions = np.zeros(self.n_levels, np.uint32)
        n = _lib.load().arcb200_grid_geometry(self.n_levels, self.log2_hashmap_size, self.base_resolution,
                                              self.per_level_scale, self.offsets.ctypes.data, self.scales.ctypes.data,
                                              self.resolutions.ctypes.data)
        if n < 0:
            raise RuntimeError('arcb200_grid_geometry failed: %s' % _lib.last_error())
        self.n_entries = int(n)
        self.n_params = self.n_entries * self.n_features
        self.out_dim = self.n_levels * self.n_features

    def _args(self):
        return (self.n_levels, self.n_features, self.offsets.ctypes.data, self.scales.ctypes.data,
                self.resolutions.ctypes.data)


def cast_f16(src, dst=None):
    """fp32 -> fp16 copy of a flat parameter tensor."""
    _cuda(src, 'src')
    if dst is None:
        dst = torch.empty(src.shape, dtype=torch.float16, device=src.device)
    _lib.call('arcb200_cast_f32_to_f16', src.data_ptr(), dst.data_ptr(), src.numel(), _stream())
    return dst


def grid_forward(xyz, table_half, geom, want_dydx=False, n_dev=None, out=None):
    """xyz (N,3) fp32 in [0,1]; table_half flat fp16 -> (N, L*F) fp16 [, dy_dx (N, L*F, 3) fp32]"""
    _cuda(xyz, 'xyz'), _cuda(table_half, 'table')
    if xyz.dtype != torch.float32 or xyz.dim() != 2 or xyz.shape[1] != 3:
        raise RuntimeError('xyz must be a (N,3) float32 tensor')
    if table_half.dtype != torch.float16 or table_half.numel() != geom.n_params:
        raise RuntimeError('table must be float16 with %d elements' % geom.n_params)
    n = xyz.shape[0]
    if out is None:
        out = torch.empty((n, geom.out_dim), dtype=torch.float16, device=xyz.device)
    dydx = torch.empty((n, geom.out_dim, 3), dtype=torch.float32, device=xyz.device) if want_dydx else None
    nl, nf, off, sc, rs = geom._args()
    _lib.call('arcb200_grid_fwd', xyz.data_ptr(), table_half.data_ptr(), nl, nf, off, sc, rs, out.data_ptr(),
              _lib.ptr(dydx), n, _lib.ptr(n_dev), _stream())
    return (out, dydx) if want_dydx else out


def grid_backward(xyz, dL_dy, geom, grad_table=None, gscale=1.0, n_dev=None):
    """Accumulates the table gradient (flat fp32, geom.n_params) of dL_dy (N, L*F) fp16|fp32."""
    _cuda(xyz, 'xyz'), _cuda(dL_dy, 'dL_dy')
    if dL_dy.dtype not in (torch.float16, torch.float32) or dL_dy.shape != (xyz.shape[0], geom.out_dim):
        raise RuntimeError('dL_dy must be (N, %d) float16/float32' % geom.out_dim)
    if grad_table is None:
        grad_table = torch.zeros(geom.n_params, dtype=torch.float32, device=xyz.device)
    nl, nf, off, sc, rs = geom._args()
    _lib.call('arcb200_grid_bwd', xyz.data_ptr(), dL_dy.data_ptr(), int(dL_dy.dtype == torch.float16), float(gscale),
              nl, nf, off, sc, rs, grad_table.data_ptr(), xyz.shape[0], _lib.ptr(n_dev), _stream())
    return grad_table


def grid_backward_input(dL_dy, dy_dx, geom, gscale=1.0):
    _cuda(dL_dy, 'dL_dy'), _cuda(dy_dx, 'dy_dx')
    n = dL_dy.shape[0]
    out = torch.empty((n, 3), dtype=torch.float32, device=dL_dy.device)
    _lib.call('arcb200_grid_bwd_input', dL_dy.data_ptr(), int(dL_dy.dtype == torch.float16), float(gscale),
              dy_dx.data_ptr(), geom.n_levels, geom.n_features, out.data_ptr(), n, _stream())
    return out


def grid_double_backward(xyz, table_half, dL_dy, v, geom, grad_table=None, want_dy=True, want_dx=True, gscale=1.0):
    """v = dL2/d(dL/dx) (N,3) -> (grad_table accumulated | None, d_dLdy (N,L*F) | None, d_x (N,3) | None)"""
    _cuda(xyz, 'xyz'), _cuda(table_half, 'table'), _cuda(dL_dy, 'dL_dy'), _cuda(v, 'v')
    n = xyz.shape[0]
    d_dy = torch.empty((n, geom.out_dim), dtype=torch.float32, device=xyz.device) if want_dy else None
    d_x = torch.empty((n, 3), dtype=torch.float32, device=xyz.device) if want_dx else None
    nl, nf, off, sc, rs = geom._args()
    _lib.call('arcb200_grid_bwd_bwd_input', xyz.data_ptr(), table_half.data_ptr(), dL_dy.data_ptr(),
              int(dL_dy.dtype == torch.float16), float(gscale), v.data_ptr(), nl, nf, off, sc, rs,
              _lib.ptr(grad_table), _lib.ptr(d_dy), _lib.ptr(d_x), n, _stream())
    return grad_table, d_dy, d_x


# ---------------------------------------------------------------------------------------------- SH
def sh_forward(dirs, degree=4, index=None, remap=True, out_dtype=torch.float16, out=None, out_col=0, n_dev=None):
    """dirs (M,3) fp32 -> (N, degree^2); index (N,) int32 selects the row of dirs per output row.
    With `out` (N, S) given, writes columns [out_col, out_col+degree^2) of it (fused concatenation)."""
    _cuda(dirs, 'dirs')
    n = dirs.shape[0] if index is None else index.shape[0]
    nd = degree * degree
    if out is None:
        out = torch.empty((n, nd), dtype=out_dtype, device=dirs.device)
        out_col = 0
    _cuda(out, 'out')
    if out.dtype not in (torch.float16, torch.float32) or out.shape[0] != n or out_col + nd > out.shape[1]:
        raise RuntimeError('sh_forward: bad output tensor')
    esz = out.element_size()
    _lib.call('arcb200_sh_fwd', dirs.data_ptr(), _lib.ptr(index), int(degree), int(bool(remap)),
              out.data_ptr() + out_col * esz, int(out.dtype == torch.float16), out.stride(0), n, _lib.ptr(n_dev), _stream())
    return out


# ---------------------------------------------------------------------------------------------- fused MLP
class MlpSpec:
    """Shape of a tcnn-style FullyFusedMLP (no biases): in_dim -> width x n_hidden -> out_dim."""

    def __init__(self, in_dim, width, n_hidden, out_dim, hidden_act='relu', out_act='none'):
        self.in_dim, self.width, self.n_hidden, self.out_dim = int(in_dim), int(width), int(n_hidden), int(out_dim)
        self.hidden_act, self.out_act = ACT_CODES[hidden_act], ACT_CODES[out_act]
        lib = _lib.load()
        self.n_params = lib.arcb200_mlp_n_params(self.in_dim, self.width, self.n_hidden, self.out_dim)
        if self.n_params < 0:
            raise RuntimeError('unsupported MLP shape: %s' % _lib.last_error())
        self.blob_bytes = lib.arcb200_mlp_blob_bytes(self.in_dim, self.width, self.n_hidden, self.out_dim)
        self.in_pad = (self.in_dim + 15) // 16 * 16
        self.out_pad = (self.out_dim + 15) // 16 * 16

    def dims(self):
        return self.in_dim, self.width, self.n_hidden, self.out_dim


def mlp_pack(params, spec, blob=None):
    """flat fp32 parameters (tcnn layout) -> fp16 tensor-core operand blob"""
    _cuda(params, 'params')
    if params.dtype != torch.float32 or params.numel() != spec.n_params:
        raise RuntimeError('params must be float32 with %d elements' % spec.n_params)
    if blob is None:
        blob = torch.empty(spec.blob_bytes, dtype=torch.uint8, device=params.device)
    _lib.call('arcb200_mlp_pack_weights', params.data_ptr(), blob.data_ptr(), *spec.dims(), _stream())
    return blob


def mlp_act_save_buffer(n, spec, device):
    """scratch for the hidden activations of n rows (forward -> backward hand-over)"""
    nbytes = _lib.load().arcb200_mlp_act_save_bytes(int(n), spec.width, spec.n_hidden)
    return torch.empty(int(nbytes), dtype=torch.uint8, device=device)


def mlp_forward(x, blob, spec, out_dtype=torch.float32, out=None, out_col=0, n_dev=None, act_save=None, exp0_out=None,
                sh=None):
    """sh = (dirs (n_dirs,3) fp32, index (N,) int32 | None, column, remap): also write the degree-4 SH encoding of each row's
    direction into `out[:, column:column+16]` (fp16 `out` only) - the radiance head's `[feat, sh]` input in one launch; an
    optional fifth element is a 1-element float tensor that the launch clears (the step's loss accumulator)"""
    _cuda(x, 'x'), _cuda(blob, 'blob')
    if x.dtype not in (torch.float16, torch.float32) or x.dim() != 2 or x.shape[1] < spec.in_dim or x.stride(1) != 1:
        raise RuntimeError('x must be (N, >=%d) float16/float32 with unit column stride' % spec.in_dim)
    n = x.shape[0]
    if out is None:
        out = torch.empty((n, spec.out_dim), dtype=out_dtype, device=x.device)
    if sh is not None:
        dirs, index, col, remap = sh[:4]
        zero_scalar = sh[4] if len(sh) > 4 else None
        if out.dtype != torch.float16 or out_col != 0:
            raise RuntimeError('the fused SH encoding needs an fp16 output written from column 0')
        _lib.call('arcb200_mlp_fwd_sh', x.data_ptr(), int(x.dtype == torch.float16), x.stride(0), blob.data_ptr(), *spec.dims(),
                  spec.hidden_act, spec.out_act, out.data_ptr(), out.stride(0), _lib.ptr(act_save), _lib.ptr(exp0_out),
                  dirs.data_ptr(), _lib.ptr(index), int(col), int(bool(remap)), _lib.ptr(zero_scalar), n, _lib.ptr(n_dev), _stream())
        return out
    _lib.call('arcb200_mlp_fwd_ex', x.data_ptr(), int(x.dtype == torch.float16), x.stride(0), blob.data_ptr(), *spec.dims(),
              spec.hidden_act, spec.out_act, out.data_ptr() + out_col * out.element_size(),
              int(out.dtype == torch.float16), out.stride(0), _lib.ptr(act_save), _lib.ptr(exp0_out), n, _lib.ptr(n_dev),
              _stream())
    return out


def mlp_backward(x, blob, spec, dy, dW=None, want_dx=True, dx_dtype=torch.float32, loss_scale=128.0, n_dev=None,
                 dx=None, act_saved=None, y_saved=None, g0=None, x0=None, clip=15.0):
    """-> (dX (N,in_dim) | None, dW flat fp32 accumulated).  act_saved / y_saved: hand-over from mlp_forward(act_save=..)
    (no forward re-computation); g0 (N,) with x0 (N-row strided view, fp16/fp32): extra gradient on output channel 0
    through exp(clamp(x0))."""
    _cuda(x, 'x'), _cuda(blob, 'blob'), _cuda(dy, 'dy')
    if dy.dtype != torch.float32 or dy.shape[0] != x.shape[0] or dy.shape[1] < spec.out_dim or dy.stride(1) != 1:
        raise RuntimeError('dy must be (N, >=%d) float32 with unit column stride' % spec.out_dim)
    n = x.shape[0]
    if dW is None:
        dW = torch.zeros(spec.n_params, dtype=torch.float32, device=x.device)
    if dx is None and want_dx:
        dx = torch.empty((n, spec.in_dim), dtype=dx_dtype, device=x.device)
    _lib.call('arcb200_mlp_bwd_ex', x.data_ptr(), int(x.dtype == torch.float16), x.stride(0), blob.data_ptr(), *spec.dims(),
              spec.hidden_act, spec.out_act, dy.data_ptr(), dy.stride(0), float(loss_scale), _lib.ptr(dx),
              int(dx is not None and dx.dtype == torch.float16), dx.stride(0) if dx is not None else 0, dW.data_ptr(),
              _lib.ptr(act_saved), _lib.ptr(y_saved), int(y_saved is not None and y_saved.dtype == torch.float16),
              y_saved.stride(0) if y_saved is not None else 0, _lib.ptr(g0), _lib.ptr(x0),
              int(x0 is not None and x0.dtype == torch.float16), x0.stride(0) if x0 is not None else 0, float(clip), n,
              _lib.ptr(n_dev), _stream())
    return dx, dW


# ---------------------------------------------------------------------------------------------- compositing
def composite_forward(geo, rad, t, offsets=None, dense_p=0, pmax=None, is_alpha=False, noise=None, add_inf_z=False,
                      white_bkg=False, bkg=None, n_rays=None, want_weights=False, want_trans=False, want_alpha=False):
    """Segmented alpha compositing.  Dense: geo/t (R,P), rad (R,P,3).  Packed: flat geo/t (N,), rad (N,3) with
    offsets (R+1,) int32 and pmax (device int32 tensor, 1 element).  Returns dict(rgb, depth, mask[, weights, ...])."""
    _cuda(geo, 'geo'), _cuda(t, 't')
    dev = geo.device
    if offsets is None:
        n_rays, dense_p = geo.shape[0], geo.shape[1]
    else:
        n_rays = offsets.shape[0] - 1
    cap = geo.numel()
    rgb = torch.empty((n_rays, 3), dtype=torch.float32, device=dev) if rad is not None else None
    depth = torch.empty((n_rays,), dtype=torch.float32, device=dev)
    acc = torch.empty((n_rays,), dtype=torch.float32, device=dev)
    w = torch.empty(geo.shape, dtype=torch.float32, device=dev) if want_weights else None
    tr = torch.empty(geo.shape, dtype=torch.float32, device=dev) if want_trans else None
    al = torch.empty(geo.shape, dtype=torch.float32, device=dev) if want_alpha else None
    bkg_per_ray = int(bkg is not None and bkg.shape[0] == n_rays and n_rays > 1)
    _lib.call('arcb200_composite_fwd', _lib.ptr(offsets), int(dense_p), cap, _lib.ptr(pmax), geo.data_ptr(),
              int(bool(is_alpha)), _lib.ptr(rad), t.data_ptr(), _lib.ptr(noise), int(bool(add_inf_z)),
              int(bool(white_bkg)), _lib.ptr(bkg), bkg_per_ray, _lib.ptr(rgb), depth.data_ptr(), acc.data_ptr(),
              _lib.ptr(w), _lib.ptr(tr), _lib.ptr(al), n_rays, _stream())
    return dict(rgb=rgb, depth=depth, mask=acc, weights=w, trans_shift=tr, alpha=al)


def composite_backward(geo, rad, t, g_rgb=None, g_depth=None, g_mask=None, g_weights=None, offsets=None, dense_p=0,
                       pmax=None, is_alpha=False, noise=None, add_inf_z=False, white_bkg=False, bkg=None,
                       max_samples_per_ray=1024):
    """-> (d_geo like geo, d_rad like rad | None)"""
    dev = geo.device
    if offsets is None:
        n_rays, dense_p = geo.shape[0], geo.shape[1]
    else:
        n_rays = offsets.shape[0] - 1
    d_geo = torch.zeros_like(geo) if offsets is not None else torch.empty_like(geo)
    d_rad = (torch.zeros_like(rad) if offsets is not None else torch.empty_like(rad)) if rad is not None else None
    bkg_per_ray = int(bkg is not None and bkg.shape[0] == n_rays and n_rays > 1)
    _lib.call('arcb200_composite_bwd', _lib.ptr(offsets), int(dense_p), geo.numel(), _lib.ptr(pmax), geo.data_ptr(),
              int(bool(is_alpha)), _lib.ptr(rad), t.data_ptr(), _lib.ptr(noise), int(bool(add_inf_z)),
              int(bool(white_bkg)), _lib.ptr(bkg), bkg_per_ray, _lib.ptr(g_rgb), _lib.ptr(g_depth), _lib.ptr(g_mask),
              _lib.ptr(g_weights), d_geo.data_ptr(), _lib.ptr(d_rad), int(max_samples_per_ray), n_rays, _stream())
    return d_geo, d_rad


def sdf_to_alpha(sdf, zvals, slope, s, clip=True):
    _cuda(sdf, 'sdf'), _cuda(zvals, 'zvals'), _cuda(slope, 'slope')
    r, p = sdf.shape
    alpha = torch.empty_like(sdf)
    s_dev = s if isinstance(s, torch.Tensor) else None
    _lib.call('arcb200_sdf_to_alpha', sdf.data_ptr(), zvals.data_ptr(), slope.data_ptr(), _lib.ptr(s_dev),
              0.0 if s_dev is not None else float(s), int(bool(clip)), alpha.data_ptr(), r, p, _stream())
    return alpha


def sdf_to_alpha_backward(sdf, zvals, slope, s, g_alpha, clip=True):
    r, p = sdf.shape
    d_sdf, d_slope = torch.empty_like(sdf), torch.empty_like(slope)
    d_s = torch.zeros(1, dtype=torch.float32, device=sdf.device)
    s_dev = s if isinstance(s, torch.Tensor) else None
    _lib.call('arcb200_sdf_to_alpha_bwd', sdf.data_ptr(), zvals.data_ptr(), slope.data_ptr(), _lib.ptr(s_dev),
              0.0 if s_dev is not None else float(s), int(bool(clip)), g_alpha.data_ptr(), d_sdf.data_ptr(),
              d_slope.data_ptr(), d_s.data_ptr(), r, p, _stream())
    return d_sdf, d_slope, d_s


def sample_cdf_gather(bins, cdf, u, eps=1e-5):
    """bins, cdf (R, nb); u (R, ns) or (ns,) -> (R, ns) unsorted inverse-CDF samples"""
    r, nb = bins.shape
    ns = u.shape[-1]
    out = torch.empty((r, ns), dtype=torch.float32, device=bins.device)
    _lib.call('arcb200_sample_cdf', bins.data_ptr(), cdf.data_ptr(), u.data_ptr(), int(u.dim() == 2), float(eps),
              out.data_ptr(), r, nb, ns, _stream())
    return out


# ---------------------------------------------------------------------------------------------- step helpers
def adam_step(p, m, v, g, step, lr, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, p_half=None, grad_scale=1.0,
              zero_grad=True, step_dev=None):
    """step_dev: device int32 tensor holding the step count (takes precedence over `step`; replayable launches)"""
    _lib.call('arcb200_adam_step', p.data_ptr(), m.data_ptr(), v.data_ptr(), g.data_ptr(), _lib.ptr(p_half), p.numel(),
              int(step), float(lr), float(beta1), float(beta2), float(eps), float(weight_decay), float(grad_scale),
              int(bool(zero_grad)), _lib.ptr(step_dev), _stream())


class MlpAdamPack:
    """Prepared argument block of arcb200_mlp_adam_pack: Adam + fp16 operand re-packing of up to 4 small networks in one
    launch.  grads: per network either a tensor (local gradient) or a list of `world` raw device pointers (every rank's
    gradient of that network, through peer memory), summed in list order."""

    def __init__(self, specs, ps, ms, vs, grads, blobs):
        n = len(specs)
        self.n = n
        self.keep = (ps, ms, vs, grads, blobs)
        self.shapes = (C.c_int * (4 * n))(*[d for sp in specs for d in sp.dims()])
        arr = lambda ts: (C.c_void_p * n)(*[t.data_ptr() for t in ts])
        self.p, self.m, self.v, self.blobs = arr(ps), arr(ms), arr(vs), arr(blobs)
        ptr_lists = [[g.data_ptr()] if torch.is_tensor(g) else [int(x) for x in g] for g in grads]
        self.world = len(ptr_lists[0])
        if any(len(x) != self.world for x in ptr_lists):
            raise RuntimeError('every network needs the same number of gradient sources')
        self.grads = (C.c_void_p * (n * self.world))(*[x for lst in ptr_lists for x in lst])

    def step(self, step, lr, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, grad_scale=1.0, zero_grad=True,
             step_dev=None):
        _lib.call('arcb200_mlp_adam_pack', self.n, self.shapes, self.p, self.m, self.v, self.grads, self.world, self.blobs,
                  int(step), float(lr), float(beta1), float(beta2), float(eps), float(weight_decay), float(grad_scale),
                  int(bool(zero_grad)), _lib.ptr(step_dev), _stream())


def huber_loss(rgb, target, delta=0.1, weight=3000.0, loss_sum=None, want_grad=True):
    """-> (loss_sum (1,) accumulated, g_rgb (R,3) | None)"""
    if loss_sum is None:
        loss_sum = torch.zeros(1, dtype=torch.float32, device=rgb.device)
    g = torch.empty_like(rgb) if want_grad else None
    _lib.call('arcb200_huber_loss', rgb.data_ptr(), target.data_ptr(), float(delta), float(weight), _lib.ptr(g),
              loss_sum.data_ptr(), rgb.shape[0], _stream())
    return loss_sum, g
