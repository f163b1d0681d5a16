"""GPU: hash-grid encoding, SH and the tcgen05 fused MLP through the C ABI against the CPU oracle (oracle/fields.py).

Stated tolerances:
  encodings  : <= 1 fp16 ulp of the fp32-blend oracle (|d| <= 2^-10 * max|feature|), and within 8 half-ulps of the
               reference's corner-by-corner fp16 accumulation;
  grid grads : fp32 atomics vs float64 oracle, rel 1e-4 of the largest entry (order of summation only);
  MLP fwd    : fp16 activations, fp32 accumulate vs float64 oracle: abs 2e-3 + rel 2e-3 (one fp16 rounding flip of a
               hidden unit moves an output by ~1e-3);
  MLP grads  : rel 1e-2 of the largest entry (fp16 gradient tiles with loss scale 128, as in the reference).
"""
import numpy as np
import pytest
import torch

from arcnerf_b200 import functional as F
from oracle import fields

pytestmark = pytest.mark.gpu


def T(a, dev, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    return t if dtype is None else t.to(dtype)


GRID_CFGS = {
    'ngp': dict(n_levels=16, n_features=2, log2_hashmap_size=19, base_resolution=16, per_level_scale=None),
    'small': dict(n_levels=6, n_features=2, log2_hashmap_size=10, base_resolution=4, per_level_scale=1.7),
    'f4': dict(n_levels=4, n_features=4, log2_hashmap_size=12, base_resolution=8, per_level_scale=2.0),
    'f1': dict(n_levels=3, n_features=1, log2_hashmap_size=9, base_resolution=4, per_level_scale=1.5),
    'f8': dict(n_levels=2, n_features=8, log2_hashmap_size=11, base_resolution=6, per_level_scale=2.0),
}


def make_grid(name):
    cfg = dict(GRID_CFGS[name])
    nf = cfg.pop('n_features')
    og = fields.grid_geometry(**cfg)
    cfg['per_level_scale'] = og[4]
    g = F.GridGeometry(n_features=nf, **cfg)
    assert np.array_equal(g.offsets, og[0]) and np.array_equal(g.scales, og[1]) and np.array_equal(g.resolutions, og[2])
    return g, og, nf


@pytest.mark.parametrize('name,n', [('ngp', 5000), ('small', 1000), ('f4', 777), ('f1', 100), ('f8', 33), ('small', 1)])
def test_grid_forward(cuda, name, n):
    g, og, nf = make_grid(name)
    rng = np.random.RandomState(0)
    table = rng.uniform(-1, 1, g.n_params).astype(np.float32)
    x = rng.uniform(0, 1, (n, 3)).astype(np.float32)
    x[0] = [0.0, 1.0, 0.5]                                   # box faces: wrap-around corners of dense levels
    th = F.cast_f16(T(table, cuda))
    assert torch.equal(th.cpu(), torch.from_numpy(table).half())
    y, dydx = F.grid_forward(T(x, cuda), th, g, want_dydx=True)
    y2 = F.grid_forward(T(x, cuda), th, g)
    assert torch.equal(y, y2)
    yo, dydxo = fields.grid_encode(x, table.reshape(-1, nf), og, nf, accumulate='fp32', with_dydx=True)
    y = y.float().cpu().numpy()
    tol = 2.0 ** -10 * np.abs(table).max()
    assert np.abs(y - yo).max() <= tol
    assert (y == yo).mean() > 0.98                          # identical apart from rare 1-ulp rounding ties
    y16 = fields.grid_encode(x, table.reshape(-1, nf), og, nf, accumulate='fp16')
    assert np.abs(y - y16).max() <= 8 * 2.0 ** -11 * np.abs(table).max()
    np.testing.assert_allclose(dydx.cpu().numpy(), dydxo, rtol=1e-4, atol=1e-4 * np.abs(dydxo).max())


@pytest.mark.parametrize('name,n,half', [('ngp', 4000, True), ('small', 3000, False), ('f4', 500, True), ('f1', 200, False)])
def test_grid_backward(cuda, name, n, half):
    g, og, nf = make_grid(name)
    rng = np.random.RandomState(1)
    x = rng.uniform(0, 1, (n, 3)).astype(np.float32)
    dy = rng.normal(size=(n, g.out_dim)).astype(np.float32)
    dy_t = T(dy, cuda, torch.float16 if half else torch.float32)
    dy_used = dy_t.float().cpu().numpy()
    grad = F.grid_backward(T(x, cuda), dy_t, g, gscale=0.5)
    go = fields.grid_backward_table(x, dy_used * 0.5, og, nf).reshape(-1)
    np.testing.assert_allclose(grad.cpu().numpy(), go, rtol=1e-4, atol=1e-4 * np.abs(go).max())
    # accumulation into an existing buffer
    grad2 = F.grid_backward(T(x, cuda), dy_t, g, grad_table=grad.clone(), gscale=0.5)
    np.testing.assert_allclose(grad2.cpu().numpy(), 2 * go, rtol=1e-4, atol=2e-4 * np.abs(go).max())
    # adjointness with the forward at full size (size-independent property): <dy, enc(table)> == <grad, table>
    table = rng.uniform(-1, 1, g.n_params).astype(np.float32)
    th = F.cast_f16(T(table, cuda))
    y = F.grid_forward(T(x, cuda), th, g).double()
    lhs = float((y * T(dy_used, cuda).double() * 0.5).sum())
    rhs = float((grad.double() * th.double()).sum())
    assert abs(lhs - rhs) <= 2e-3 * abs(lhs) + 1e-2


def test_grid_input_gradient_and_double_backward(cuda):
    g, og, nf = make_grid('small')
    rng = np.random.RandomState(2)
    n = 600
    table = rng.uniform(-1, 1, g.n_params).astype(np.float32)
    x = rng.uniform(0.01, 0.99, (n, 3)).astype(np.float32)
    dy = rng.normal(size=(n, g.out_dim)).astype(np.float32)
    v = rng.normal(size=(n, 3)).astype(np.float32)
    th = F.cast_f16(T(table, cuda))
    _, dydx = F.grid_forward(T(x, cuda), th, g, want_dydx=True)
    dx = F.grid_backward_input(T(dy, cuda), dydx, g)
    _, dydxo = fields.grid_encode(x, table.reshape(-1, nf), og, nf, with_dydx=True)
    dxo = np.einsum('nk,nkd->nd', dy.astype(np.float64), dydxo.astype(np.float64))
    np.testing.assert_allclose(dx.cpu().numpy(), dxo, rtol=1e-4, atol=1e-4 * np.abs(dxo).max())
    gt, d_dy, d_x = F.grid_double_backward(T(x, cuda), th, T(dy, cuda), T(v, cuda), g,
                                           grad_table=torch.zeros(g.n_params, device=cuda))
    gto, d_dyo, d_xo = fields.grid_double_backward(x, table.reshape(-1, nf), dy, v, og, nf)
    np.testing.assert_allclose(gt.cpu().numpy(), gto.reshape(-1), rtol=1e-4, atol=1e-4 * np.abs(gto).max())
    np.testing.assert_allclose(d_dy.cpu().numpy(), d_dyo, rtol=1e-4, atol=1e-4 * np.abs(d_dyo).max())
    np.testing.assert_allclose(d_x.cpu().numpy(), d_xo, rtol=1e-4, atol=1e-4 * np.abs(d_xo).max())


@pytest.mark.parametrize('deg', [1, 2, 3, 4, 5])
def test_sh(cuda, deg):
    rng = np.random.RandomState(3)
    d = rng.normal(size=(1000, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    o32 = F.sh_forward(T(d, cuda), deg, out_dtype=torch.float32).cpu().numpy()
    ref = fields.sh_encode(d, deg, remap=True)
    np.testing.assert_allclose(o32, ref, rtol=1e-5, atol=2e-6)
    o16 = F.sh_forward(T(d, cuda), deg, out_dtype=torch.float16).float().cpu().numpy()
    np.testing.assert_allclose(o16, ref.astype(np.float16).astype(np.float32), rtol=2e-3, atol=1e-3)
    # gather by ray index + write into a wider row (the fused [feat | sh] input of the radiance net)
    idx = rng.randint(0, 1000, 333).astype(np.int32)
    buf = torch.full((333, 48), 7.0, dtype=torch.float16, device=cuda)
    F.sh_forward(T(d, cuda), deg, index=T(idx, cuda), out=buf, out_col=16)
    got = buf.float().cpu().numpy()
    assert np.all(got[:, :16] == 7.0) and np.all(got[:, 16 + deg * deg:] == 7.0)
    np.testing.assert_array_equal(got[:, 16:16 + deg * deg], o16[idx])


MLP_CFGS = [
    # (in, width, hidden, out, hidden_act, out_act) -- the first two are the NGP geometry / radiance heads
    (32, 64, 1, 16, 'relu', 'none'),
    (32, 64, 2, 3, 'relu', 'sigmoid'),
    (38, 64, 2, 3, 'relu', 'sigmoid'),      # NeuS radiance: 38 -> 48 inputs, pad columns are ones
    (16, 128, 3, 20, 'softplus', 'exp'),
    (64, 32, 1, 1, 'relu', 'none'),
]


def close_but(actual, desired, rtol, atol, bad_frac=2e-5, hard=10.0):
    """allclose, except that a fraction `bad_frac` of the entries may miss the tolerance by up to `hard` times:
    a hidden unit whose pre-activation rounds to the other side of zero (fp16) flips its ReLU gate in one row."""
    err = np.abs(actual - desired)
    lim = atol + rtol * np.abs(desired)
    bad = err > lim
    assert bad.mean() <= bad_frac, 'too many entries out of tolerance: %g' % bad.mean()
    assert np.all(err <= hard * lim), 'entry far out of tolerance: %g' % (err / lim).max()


def rand_params(rng, n, scale=0.35, cfg=None):
    """uniform weights; with cfg, xavier-uniform per matrix like the reference's initialiser"""
    if cfg is None:
        return rng.uniform(-scale, scale, n).astype(np.float32)
    shapes, _, _ = fields.mlp_shapes(*cfg[:4])
    return np.concatenate([rng.uniform(-1, 1, o * i) * np.sqrt(6.0 / (o + i)) for o, i in shapes]).astype(np.float32)


@pytest.mark.parametrize('cfg', MLP_CFGS)
@pytest.mark.parametrize('n', [1, 128, 1000, 20000])
def test_mlp_forward(cuda, cfg, n):
    in_dim, width, n_hidden, out_dim, ha, oa = cfg
    spec = F.MlpSpec(in_dim, width, n_hidden, out_dim, ha, oa)
    rng = np.random.RandomState(4)
    params = rand_params(rng, spec.n_params, cfg=cfg if oa == 'exp' else None)
    x = rng.uniform(-1, 1, (n, in_dim)).astype(np.float32)
    blob = F.mlp_pack(T(params, cuda), spec)
    ref = fields.mlp_forward(x, params, in_dim, width, n_hidden, out_dim, ha, oa)
    for xdt in (torch.float32, torch.float16):
        y = F.mlp_forward(T(x, cuda, xdt), blob, spec).cpu().numpy()
        np.testing.assert_allclose(y, ref, rtol=2e-3, atol=2e-3 * max(1.0, np.abs(ref).max()))
    yh = F.mlp_forward(T(x, cuda), blob, spec, out_dtype=torch.float16).float().cpu().numpy()
    np.testing.assert_allclose(yh, ref, rtol=4e-3, atol=4e-3 * max(1.0, np.abs(ref).max()))


@pytest.mark.parametrize('cfg', [c for c in MLP_CFGS if c[1] in (64, 128)])
@pytest.mark.parametrize('n', [1, 300, 20000])
def test_mlp_backward(cuda, cfg, n):
    in_dim, width, n_hidden, out_dim, ha, oa = cfg
    spec = F.MlpSpec(in_dim, width, n_hidden, out_dim, ha, oa)
    rng = np.random.RandomState(5)
    params = rand_params(rng, spec.n_params, cfg=cfg if oa == 'exp' else None)
    x = rng.uniform(-1, 1, (n, in_dim)).astype(np.float32)
    dy = (rng.normal(size=(n, out_dim)) * 1e-2).astype(np.float32)
    blob = F.mlp_pack(T(params, cuda), spec)
    dx, dw = F.mlp_backward(T(x, cuda), blob, spec, T(dy, cuda))
    dxo, dwo = fields.mlp_backward(x, params, dy, in_dim, width, n_hidden, out_dim, ha, oa)
    close_but(dx.cpu().numpy(), dxo, rtol=1e-2, atol=1e-2 * np.abs(dxo).max())
    close_but(dw.cpu().numpy(), dwo, rtol=1e-2, atol=1e-2 * np.abs(dwo).max(), bad_frac=0.0)
    # linearity in dy and accumulation into dW (size-independent properties)
    dx2, dw2 = F.mlp_backward(T(x, cuda), blob, spec, T(2 * dy, cuda), dW=dw.clone(), want_dx=True)
    np.testing.assert_allclose(dw2.cpu().numpy(), 3 * dw.cpu().numpy(), rtol=2e-2, atol=2e-2 * np.abs(dwo).max())
    _, dw3 = F.mlp_backward(T(x, cuda), blob, spec, T(dy, cuda), want_dx=False)
    np.testing.assert_allclose(dw3.cpu().numpy(), dw.cpu().numpy(), rtol=1e-3, atol=1e-3 * np.abs(dwo).max())


def test_mlp_rejects_bad_shapes(cuda):
    with pytest.raises(RuntimeError):
        F.MlpSpec(32, 60, 1, 16)            # width not a multiple of 16
    with pytest.raises(RuntimeError):
        F.MlpSpec(32, 64, 4, 16)            # too many hidden layers
    spec = F.MlpSpec(32, 32, 1, 16)
    blob = F.mlp_pack(torch.zeros(spec.n_params, device=cuda), spec)
    with pytest.raises(RuntimeError):      # training needs width 64 or 128
        F.mlp_backward(torch.zeros(4, 32, device=cuda), blob, spec, torch.zeros(4, 16, device=cuda))


def test_adam_and_huber(cuda):
    rng = np.random.RandomState(6)
    n = 100003
    p = rng.normal(size=n).astype(np.float32); g = rng.normal(size=n).astype(np.float32)
    pt = torch.tensor(p, requires_grad=True)
    opt = torch.optim.Adam([pt], lr=1e-2, betas=(0.9, 0.99), eps=1e-15, weight_decay=1e-6)
    pc, m, v = T(p, cuda), torch.zeros(n, device=cuda), torch.zeros(n, device=cuda)
    ph = torch.empty(n, dtype=torch.float16, device=cuda)
    for step in range(1, 4):
        gs = (g * step).astype(np.float32)
        pt.grad = torch.tensor(gs)
        opt.step()
        gc = T(gs, cuda)
        F.adam_step(pc, m, v, gc, step, 1e-2, 0.9, 0.99, 1e-15, 1e-6, p_half=ph)
        assert float(gc.abs().max()) == 0.0                       # gradient cleared for the next step
    np.testing.assert_allclose(pc.cpu().numpy(), pt.detach().numpy(), rtol=1e-5, atol=1e-6)
    assert torch.equal(ph, pc.half())
    from oracle import render
    rgb = rng.uniform(0, 1, (4096, 3)).astype(np.float32); tgt = rng.uniform(0, 1, (4096, 3)).astype(np.float32)
    rt = torch.tensor(rgb, requires_grad=True)
    lo = render.huber_loss(rt, torch.tensor(tgt)); lo.backward()
    ls, gr = F.huber_loss(T(rgb, cuda), T(tgt, cuda))
    np.testing.assert_allclose(float(ls), float(lo.detach()), rtol=1e-5)
    np.testing.assert_allclose(gr.cpu().numpy(), rt.grad.numpy(), rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize('cfg', [MLP_CFGS[0], MLP_CFGS[1], MLP_CFGS[2]])
@pytest.mark.parametrize('n', [1, 300, 20000])
def test_mlp_saved_activations_and_density_extras(cuda, cfg, n):
    """The forward's activation hand-over + the backward that skips re-computation give the re-computing backward's
    results; exp0_out / (g0, x0) reproduce TruncExp forward / backward on output channel 0."""
    in_dim, width, n_hidden, out_dim, ha, oa = cfg
    spec = F.MlpSpec(in_dim, width, n_hidden, out_dim, ha, oa)
    rng = np.random.RandomState(7)
    params = rand_params(rng, spec.n_params)
    x = T(rng.uniform(-1, 1, (n, in_dim)).astype(np.float32), cuda, torch.float16)
    dy = T((rng.normal(size=(n, out_dim)) * 1e-2).astype(np.float32), cuda)
    blob = F.mlp_pack(T(params, cuda), spec)
    save = F.mlp_act_save_buffer(n, spec, cuda)
    exp0 = torch.empty(n, device=cuda)
    y16 = torch.empty((n, out_dim), dtype=torch.float16, device=cuda)
    F.mlp_forward(x, blob, spec, out=y16, act_save=save, exp0_out=exp0)
    y_plain = F.mlp_forward(x, blob, spec, out_dtype=torch.float16)
    assert torch.equal(y16, y_plain)
    np.testing.assert_allclose(exp0.cpu().numpy(), np.exp(y16[:, 0].float().cpu().numpy()), rtol=1e-6)
    dx0, dw0 = F.mlp_backward(x, blob, spec, dy)
    dx1, dw1 = F.mlp_backward(x, blob, spec, dy, act_saved=save, y_saved=y16)
    np.testing.assert_allclose(dx1.cpu().numpy(), dx0.cpu().numpy(), rtol=2e-3, atol=2e-3 * float(dx0.abs().max()))
    np.testing.assert_allclose(dw1.cpu().numpy(), dw0.cpu().numpy(), rtol=2e-3, atol=2e-3 * float(dw0.abs().max()))
    # extra gradient on channel 0 through exp(clamp(x0))
    g0 = T((rng.normal(size=n) * 1e-2).astype(np.float32), cuda)
    dy2 = dy.clone()
    dy2[:, 0] += g0 * torch.exp(y16[:, 0].float().clamp(-15, 15))
    dx2, dw2 = F.mlp_backward(x, blob, spec, dy2, act_saved=save, y_saved=y16)
    dx3, dw3 = F.mlp_backward(x, blob, spec, dy, act_saved=save, y_saved=y16, g0=g0, x0=y16, clip=15.0)
    np.testing.assert_allclose(dx3.cpu().numpy(), dx2.cpu().numpy(), rtol=1e-3, atol=1e-3 * float(dx2.abs().max()))
    np.testing.assert_allclose(dw3.cpu().numpy(), dw2.cpu().numpy(), rtol=1e-3, atol=1e-3 * float(dw2.abs().max()))
    if oa != 'none':
        with pytest.raises(RuntimeError):
            F.mlp_backward(x, blob, spec, dy, act_saved=save)        # saved path needs the saved outputs


def test_mlp_forward_with_fused_sh_equals_separate_launches(cuda):
    """geometry head + SH of the per-row direction in one launch == mlp_forward followed by sh_forward"""
    spec = F.MlpSpec(32, 64, 1, 16, 'relu', 'none')
    rng = np.random.RandomState(8)
    n, n_rays = 5000, 300
    params = rand_params(rng, spec.n_params)
    blob = F.mlp_pack(T(params, cuda), spec)
    x = T(rng.uniform(-1, 1, (n, 32)).astype(np.float32), cuda, torch.float16)
    dirs = rng.normal(size=(n_rays, 3)).astype(np.float32)
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    dirs_d = T(dirs, cuda)
    idx = T(rng.randint(0, n_rays, n).astype(np.int32), cuda)
    a = torch.zeros((n, 32), dtype=torch.float16, device=cuda)
    sig_a = torch.empty(n, device=cuda)
    F.mlp_forward(x, blob, spec, out=a, exp0_out=sig_a)
    F.sh_forward(dirs_d, 4, index=idx, remap=True, out=a, out_col=16)
    b = torch.zeros((n, 32), dtype=torch.float16, device=cuda)
    sig_b = torch.empty(n, device=cuda)
    F.mlp_forward(x, blob, spec, out=b, exp0_out=sig_b, sh=(dirs_d, idx, 16, True))
    assert torch.equal(a, b) and torch.equal(sig_a, sig_b)
    assert float(b[:, 16:].float().abs().max()) > 0.1
    with pytest.raises(RuntimeError):
        F.mlp_forward(x, blob, spec, out=torch.zeros((n, 32), device=cuda), sh=(dirs_d, idx, 16, True))     # fp32 output
