"""CPU oracle for the learned-field part of the hot path: multiresolution hash-grid encoding, spherical
harmonics and the small fused MLP.  TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

These algorithms live in tiny-cuda-nn, vendored as a submodule of the reference at
/root/reference/include/tiny-cuda-nn (commit 8e6e242f36dd197134c9b9275a8e5108a8e3af78).  Its kernels are
__device__-only (they cannot run in a CPU container), so this is a numpy restatement of the published
arithmetic, each function citing the lines it follows:
    encodings/grid.h:111-204        hash / dense index, level scale and resolution
    encodings/grid.h:220-392        kernel_grid (forward, dy_dx)
    encodings/grid.h:394-517        kernel_grid_backward
    encodings/grid.h:549-895        input gradient and its double backward
    encodings/grid.h:982-1012       offset table
    common_device.h:402-443         pos_fract
    encodings/spherical_harmonics.h:46-152   kernel_sh
    src/fully_fused_mlp.cu:315-557,658-671   layer order, weight layout, no biases, fp16 storage
    encodings/identity.h:46-69      input padding with ones
PARITY PINNING: pinned against tiny-cuda-nn itself.  The reference's own tiny-cuda-nn sources are built for sm_100 by
`make -C oracle tcnn` (oracle/tcnn_harness.cu -> oracle/_ref/libtcnnref.so), executed on a B200 by
tests/golden/make_golden_tcnn.py, and the outputs are committed as tests/golden/tcnn_golden.npz; tests/test_oracle_tcnn.py
checks every function below against that file (grid forward in fp16-accumulate mode is BIT-EXACT with the level scales the
device computed; gradients, double backward, SH and the MLPs within the tolerances stated there), and
tests/test_tcnn_parity_gpu.py checks the CUDA path against the same file.  The SH basis is additionally checked against
ArcNerf's torch SHEmbedder (sh_encoder.py:126-199) and the trilinear weights against Volume (tests/golden/make_golden_torch.py).
"""
import numpy as np

PRIME_Y = np.uint64(2654435761)
PRIME_Z = np.uint64(805459861)
U32 = np.uint64(0xFFFFFFFF)


def grid_geometry(n_levels=16, log2_hashmap_size=19, base_resolution=16, per_level_scale=None, max_res=2048):
    """-> offsets (L+1,) uint32 [entries], scales (L,) float32, resolutions (L,) uint32, hashed (L,) bool.
    grid.h:195-204 (scale/resolution) and :982-1012 (offsets: dense size rounded up to 8, capped at 2^T)."""
    if per_level_scale is None:
        # ArcNerf: exp(log(max_res/base_res)/(L-1)) evaluated in float32 torch (hashgrid_encoder.py:81)
        per_level_scale = float(np.exp(np.log(np.float32(max_res / base_resolution)) / np.float32(n_levels - 1)))
    import ctypes as C
    from . import sampler
    lib = sampler.Sampler('oracle').lib
    lib.orc_grid_geometry.restype = C.c_int64
    offsets = np.zeros(n_levels + 1, np.uint32)
    scales = np.zeros(n_levels, np.float32)
    res = np.zeros(n_levels, np.uint32)
    lib.orc_grid_geometry(int(n_levels), int(log2_hashmap_size), int(base_resolution), C.c_float(per_level_scale),
                          offsets.ctypes.data_as(C.c_void_p), scales.ctypes.data_as(C.c_void_p),
                          res.ctypes.data_as(C.c_void_p))
    hashed = np.zeros(n_levels, bool)
    for l in range(n_levels):
        params, r, stride = int(offsets[l + 1] - offsets[l]), int(res[l]), 1
        for _ in range(3):                      # grid.h:170-187: dense while the stride fits
            if stride <= params:
                stride *= r
        hashed[l] = params < stride
    return offsets, scales, res, hashed, float(per_level_scale)


def _locate(x, scale):
    """pos = fma(x, scale, 0.5); cell = floor(pos); frac = pos - cell   (common_device.h:402-443).
    The device contracts the multiply-add; emulate the single rounding through float64."""
    pos = (x.astype(np.float64) * np.float64(scale) + 0.5).astype(np.float32)
    fl = np.floor(pos)
    return fl.astype(np.int64).astype(np.uint64) & U32, (pos - fl).astype(np.float32)


def _index(gx, gy, gz, res, size, hashed):
    if hashed:
        idx = (gx ^ ((gy * PRIME_Y) & U32) ^ ((gz * PRIME_Z) & U32)) & U32
    else:
        r = np.uint64(res)
        idx = (gx + ((gy * r) & U32) + ((gz * ((r * r) & U32)) & U32)) & U32
    return (idx % np.uint64(size)).astype(np.int64)


def _corners(x, geom, level):
    offsets, scales, res, hashed = geom[:4]
    size = int(offsets[level + 1] - offsets[level])
    g0, f0 = _locate(x[:, 0], scales[level])
    g1, f1 = _locate(x[:, 1], scales[level])
    g2, f2 = _locate(x[:, 2], scales[level])
    out = []
    for k in range(8):
        cx, cy, cz = k & 1, (k >> 1) & 1, (k >> 2) & 1
        idx = _index((g0 + np.uint64(cx)) & U32, (g1 + np.uint64(cy)) & U32, (g2 + np.uint64(cz)) & U32,
                     int(res[level]), size, bool(hashed[level]))
        out.append((idx + int(offsets[level]), (cx, cy, cz)))
    return out, (f0, f1, f2)


def grid_encode(x, table, geom, n_features=2, accumulate='fp32', with_dydx=False):
    """x (N,3) float32 in [0,1]; table (entries, F) - values are rounded to fp16 first (the reference's parameter
    precision).  accumulate='fp32': blend in fp32, round the result to fp16 once (what libarcb200 does);
    accumulate='fp16': round to fp16 after every corner like kernel_grid (grid.h:339-341).
    -> (N, L*F) float32 holding fp16-representable values [, dy_dx (N, L*F, 3) float32]."""
    x = np.ascontiguousarray(x, np.float32)
    tab = np.asarray(table, np.float32).astype(np.float16).astype(np.float32).reshape(-1, n_features)
    L = len(geom[1])
    n = x.shape[0]
    out = np.zeros((n, L * n_features), np.float32)
    dydx = np.zeros((n, L * n_features, 3), np.float32) if with_dydx else None
    one = np.float32(1.0)
    for l in range(L):
        corners, (f0, f1, f2) = _corners(x, geom, l)
        w0, w1, w2 = (one - f0, f0), (one - f1, f1), (one - f2, f2)
        if accumulate == 'fp16':
            acc = np.zeros((n, n_features), np.float16)
        else:
            acc = np.zeros((n, n_features), np.float32)
        g = np.zeros((n, n_features, 3), np.float64)
        for idx, (cx, cy, cz) in corners:
            w = (w0[cx] * w1[cy]) * w2[cz]
            v = tab[idx]
            if accumulate == 'fp16':
                term = (w[:, None] * v).astype(np.float16)
                acc = (acc.astype(np.float64) + term.astype(np.float64)).astype(np.float16)
            else:
                acc = acc + w[:, None] * v
            if with_dydx:
                g[:, :, 0] += ((1.0 if cx else -1.0) * (w1[cy] * w2[cz]))[:, None] * v
                g[:, :, 1] += ((1.0 if cy else -1.0) * (w0[cx] * w2[cz]))[:, None] * v
                g[:, :, 2] += ((1.0 if cz else -1.0) * (w0[cx] * w1[cy]))[:, None] * v
        out[:, l * n_features:(l + 1) * n_features] = acc.astype(np.float16).astype(np.float32)
        if with_dydx:
            dydx[:, l * n_features:(l + 1) * n_features, :] = (g * np.float64(geom[1][l])).astype(np.float32)
    return (out, dydx) if with_dydx else out


def grid_backward_table(x, dL_dy, geom, n_features=2):
    """Gradient w.r.t. the table (entries, F), float64 accumulation (kernel_grid_backward, grid.h:394-517)."""
    x = np.ascontiguousarray(x, np.float32)
    L = len(geom[1])
    grad = np.zeros((int(geom[0][-1]), n_features), np.float64)
    one = np.float32(1.0)
    for l in range(L):
        corners, (f0, f1, f2) = _corners(x, geom, l)
        w0, w1, w2 = (one - f0, f0), (one - f1, f1), (one - f2, f2)
        gy = np.asarray(dL_dy[:, l * n_features:(l + 1) * n_features], np.float64)
        for idx, (cx, cy, cz) in corners:
            w = ((w0[cx] * w1[cy]) * w2[cz]).astype(np.float64)
            np.add.at(grad, idx, gy * w[:, None])
    return grad


def grid_double_backward(x, table, dL_dy, v, geom, n_features=2):
    """Double backward of dL/dx = dL/dy . dy/dx w.r.t. (table, dL/dy, x) given v = dL2/d(dL/dx) (N,3), float64.
    grid.h:578-699 (table), :870-895 (dL/dy), :701-868 (x)."""
    x = np.ascontiguousarray(x, np.float32)
    tab = np.asarray(table, np.float32).astype(np.float16).astype(np.float64).reshape(-1, n_features)
    L = len(geom[1])
    n = x.shape[0]
    g_tab = np.zeros_like(tab)
    d_dy = np.zeros((n, L * n_features), np.float64)
    d_x = np.zeros((n, 3), np.float64)
    v = np.asarray(v, np.float64)
    for l in range(L):
        corners, f = _corners(x, geom, l)
        f = [fi.astype(np.float64) for fi in f]
        w = [(1.0 - fi, fi) for fi in f]
        s = np.float64(geom[1][l])
        gy = np.asarray(dL_dy[:, l * n_features:(l + 1) * n_features], np.float64)
        for idx, c in corners:
            sg = [1.0 if ci else -1.0 for ci in c]
            dw = [sg[0] * w[1][c[1]] * w[2][c[2]], sg[1] * w[0][c[0]] * w[2][c[2]], sg[2] * w[0][c[0]] * w[1][c[1]]]
            coef = s * (v[:, 0] * dw[0] + v[:, 1] * dw[1] + v[:, 2] * dw[2])
            np.add.at(g_tab, idx, gy * coef[:, None])
            tv = tab[idx]
            d_dy[:, l * n_features:(l + 1) * n_features] += coef[:, None] * tv
            h = {(0, 1): sg[0] * sg[1] * w[2][c[2]], (0, 2): sg[0] * sg[2] * w[1][c[1]], (1, 2): sg[1] * sg[2] * w[0][c[0]]}
            gt = (gy * tv).sum(1) * s * s
            d_x[:, 0] += gt * (v[:, 1] * h[(0, 1)] + v[:, 2] * h[(0, 2)])
            d_x[:, 1] += gt * (v[:, 0] * h[(0, 1)] + v[:, 2] * h[(1, 2)])
            d_x[:, 2] += gt * (v[:, 0] * h[(0, 2)] + v[:, 1] * h[(1, 2)])
    return g_tab, d_dy, d_x


def sh_encode(dirs, degree=4, remap=True):
    """Real SH basis of kernel_sh (spherical_harmonics.h:46-152), fp32.  remap: apply ArcNerf's (d+1)/2 first
    (sh_encoder.py:116), then the encoding's 2x-1."""
    d = np.asarray(dirs, np.float32)
    if remap:
        d = (d + np.float32(1.0)) / np.float32(2.0)
    d = (d.astype(np.float64) * 2.0 - 1.0).astype(np.float32)
    x, y, z = d[:, 0], d[:, 1], d[:, 2]
    xy, xz, yz, x2, y2, z2 = x * y, x * z, y * z, x * x, y * y, z * z
    o = [np.full_like(x, 0.28209479177387814)]
    if degree > 1:
        o += [-0.48860251190291987 * y, 0.48860251190291987 * z, -0.48860251190291987 * x]
    if degree > 2:
        o += [1.0925484305920792 * xy, -1.0925484305920792 * yz, 0.94617469575755997 * z2 - 0.31539156525251999,
              -1.0925484305920792 * xz, 0.54627421529603959 * x2 - 0.54627421529603959 * y2]
    if degree > 3:
        o += [0.59004358992664352 * y * (-3.0 * x2 + y2), 2.8906114426405538 * xy * z,
              0.45704579946446572 * y * (1.0 - 5.0 * z2), 0.3731763325901154 * z * (5.0 * z2 - 3.0),
              0.45704579946446572 * x * (1.0 - 5.0 * z2), 1.4453057213202769 * z * (x2 - y2),
              0.59004358992664352 * x * (-x2 + 3.0 * y2)]
    if degree > 4:
        x4, y4, z4 = x2 * x2, y2 * y2, z2 * z2
        o += [2.5033429417967046 * xy * (x2 - y2), 1.7701307697799304 * yz * (-3.0 * x2 + y2),
              0.94617469575756008 * xy * (7.0 * z2 - 1.0), 0.66904654355728921 * yz * (3.0 - 7.0 * z2),
              -3.1735664074561294 * z2 + 3.7024941420321507 * z4 + 0.31735664074561293,
              0.66904654355728921 * xz * (3.0 - 7.0 * z2), 0.47308734787878004 * (x2 - y2) * (7.0 * z2 - 1.0),
              1.7701307697799304 * xz * (-x2 + 3.0 * y2),
              -3.7550144126950569 * x2 * y2 + 0.62583573544917614 * x4 + 0.62583573544917614 * y4]
    return np.stack(o, -1).astype(np.float32)


# ------------------------------------------------------------------------------------------------ MLP
def _act(name, x):
    if name in (None, 'none', 'None', 0):
        return x
    if name in ('relu', 'ReLU', 1):
        return np.maximum(x, 0)
    if name in ('sigmoid', 'Sigmoid', 2):
        return 1.0 / (1.0 + np.exp(-x))
    if name in ('softplus', 'Softplus', 3):
        return np.log(np.exp(x * 10.0) + 1.0) / 10.0
    if name in ('exp', 'Exponential', 4):
        return np.exp(x)
    raise ValueError(name)


def _act_grad_from_out(name, y):
    if name in (None, 'none', 'None', 0):
        return np.ones_like(y)
    if name in ('relu', 'ReLU', 1):
        return (y > 0).astype(y.dtype)
    if name in ('sigmoid', 'Sigmoid', 2):
        return y * (1.0 - y)
    if name in ('softplus', 'Softplus', 3):
        return 1.0 - np.exp(-y * 10.0)
    if name in ('exp', 'Exponential', 4):
        return y
    raise ValueError(name)


def mlp_shapes(in_dim, width, n_hidden, out_dim):
    in_pad, out_pad = (in_dim + 15) // 16 * 16, (out_dim + 15) // 16 * 16
    shapes = [(width, in_pad)] + [(width, width)] * (n_hidden - 1) + [(out_pad, width)]
    return shapes, in_pad, out_pad


def mlp_split(params, in_dim, width, n_hidden, out_dim):
    """flat tcnn-layout parameters -> list of (out x in) matrices (fully_fused_mlp.cu:658-671)"""
    shapes, _, _ = mlp_shapes(in_dim, width, n_hidden, out_dim)
    mats, off = [], 0
    for (o, i) in shapes:
        mats.append(np.asarray(params[off:off + o * i]).reshape(o, i))
        off += o * i
    assert off == len(params)
    return mats


def mlp_forward(x, params, in_dim, width, n_hidden, out_dim, hidden_act='relu', out_act='none', keep=False):
    """fp16 weights and activations, fp32 (here float64-exact) accumulation; inputs padded with ones
    (identity.h:63-64).  Returns (n, out_dim) float32 [and the per-layer fp16 activations]."""
    mats = [m.astype(np.float16).astype(np.float64) for m in mlp_split(np.asarray(params, np.float32), in_dim, width, n_hidden, out_dim)]
    _, in_pad, _ = mlp_shapes(in_dim, width, n_hidden, out_dim)
    n = x.shape[0]
    h = np.ones((n, in_pad), np.float64)
    h[:, :in_dim] = np.asarray(x, np.float32).astype(np.float16).astype(np.float64)
    acts = [h]
    for l, w in enumerate(mats):
        z = h @ w.T
        if l < len(mats) - 1:
            h = _act(hidden_act, z).astype(np.float16).astype(np.float64)
            acts.append(h)
        else:
            y = _act(out_act, z)
    out = y[:, :out_dim].astype(np.float32)
    return (out, acts, y) if keep else out


def mlp_backward(x, params, dy, in_dim, width, n_hidden, out_dim, hidden_act='relu', out_act='none'):
    """-> dX (n, in_dim), dW flat (parameter layout), float64 math on the fp16-rounded forward."""
    mats = [m.astype(np.float16).astype(np.float64) for m in mlp_split(np.asarray(params, np.float32), in_dim, width, n_hidden, out_dim)]
    out, acts, y = mlp_forward(x, params, in_dim, width, n_hidden, out_dim, hidden_act, out_act, keep=True)
    _, _, out_pad = mlp_shapes(in_dim, width, n_hidden, out_dim)
    g = np.zeros((x.shape[0], out_pad), np.float64)
    g[:, :out_dim] = np.asarray(dy, np.float64) * _act_grad_from_out(out_act, y[:, :out_dim])
    dws = [None] * len(mats)
    for l in range(len(mats) - 1, -1, -1):
        dws[l] = g.T @ acts[l]
        g = g @ mats[l]
        if l > 0:
            g = g * _act_grad_from_out(hidden_act, acts[l])
    return g[:, :in_dim], np.concatenate([d.reshape(-1) for d in dws])
