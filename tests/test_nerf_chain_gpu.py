"""The vanilla-NeRF network pair as one tensor-core chain (csrc/dense_chain.cu, plugin/nerf_mlp.py) against the REFERENCE's
own GeoNet + RadianceNet at full width (tests/golden/nerf_full_golden.npz, made by tests/golden/make_golden_nerf_full.py from
the reference's modules with the seeded weights of tests/nerf_full_cases.py).
Tolerances: the reference computes in fp32; the chain uses fp16 operands (weights, embeddings, activations) with fp32
accumulation over 9 + 2 layers -> sigma abs 3e-2 (values of order 1), rgb abs 4e-3.  Gradients against the fp32 reference:
relative L2 error 2e-1 per tensor (measured: 1e-3 at the last layers growing to 1.3e-1 at the first) - about 0.3 % of the
ReLU gates per layer have a pre-activation within the accumulated fp16 rounding of zero and open differently than in fp32;
a flipped gate changes its gradient entry by 100 %, so every layer adds about sqrt(0.003) = 5 % of relative L2 difference.  The
kernel logic itself is checked tightly: the forward against the same network in torch with weights, embeddings and
activations rounded to fp16 exactly where the kernels round them (5e-4), the backward against manual backpropagation through
the activations the forward kernel saved, i.e. with identical gates (5e-3)."""
import os

import numpy as np
import pytest
import torch

import nerf_full_cases as nc

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'nerf_full_golden.npz'))


def build(name):
    from arcnerf_b200.plugin.nerf_mlp import FusedNeRFNets
    d_x, nfx, nfv, W, D, skips, W_rad = nc.CASES[name]
    nets = FusedNeRFNets(d_x, nfx, nfv, W, D, skips, W_rad)
    gs, rs = nc.weights(name)
    nets.load_reference_state(gs, rs)
    return nets


@pytest.mark.parametrize('name', list(nc.CASES))
def test_forward_matches_reference_modules(name):
    nets = build(name)
    x, dirs, _, _ = (torch.from_numpy(a).cuda() for a in nc.inputs(name))
    with torch.no_grad():
        sigma, rgb = nets(x, dirs)
    torch.cuda.synchronize()
    assert sigma.shape == (nc.N_PTS, 1) and rgb.shape == (nc.N_PTS, 3)
    ref_s, ref_c = G[name + '/sigma'], G[name + '/rgb']
    assert np.abs(sigma.cpu().numpy() - ref_s).max() < 3e-2 * max(1.0, np.abs(ref_s).max())
    assert np.abs(rgb.cpu().numpy() - ref_c).max() < 4e-3


@pytest.mark.parametrize('name', list(nc.CASES))
def test_parameter_gradients_match_reference_autograd(name):
    nets = build(name)
    x, dirs, w_s, w_c = (torch.from_numpy(a).cuda() for a in nc.inputs(name))
    sigma, rgb = nets(x, dirs)
    ((sigma * w_s).sum() + (rgb * w_c).sum()).backward()
    torch.cuda.synchronize()
    g = nets.flat.grad
    off = 0
    n_geo = nets.arch[4] + 1
    for li, (o, i) in enumerate(nets.shapes):
        tag, k = ('geo', li) if li < n_geo else ('rad', li - n_geo)
        gw = g[off:off + o * i].view(o, i).cpu().numpy()
        gb = g[off + o * i:off + o * i + o].cpu().numpy()
        off += o * i + o
        key = '%s/grad/%s/layers.%d' % (name, tag, k)
        scale = float(G[key + '.weight.absmax'])
        for got, ref in ((gw[:8], G[key + '.weight.rows8']), (gw.sum(0), G[key + '.weight.colsum']), (gb, G[key + '.bias'])):
            assert np.abs(got - ref).max() < 2.5e-1 * max(np.abs(ref).max(), scale), key
            assert np.linalg.norm(got - ref) < 2e-1 * np.linalg.norm(ref) + 1e-6, key


def _emulated(nets, x, dirs):
    """GeoNet + RadianceNet forward in torch with the kernels' roundings: fp16 weights, embeddings and activations, fp32 sums"""
    d_x, nfx, nfv, W, D, skip, W_rad = nets.arch
    h16 = lambda t: t.half().float()

    def embed(v, nf):
        parts = [v]
        for k in range(nf):
            parts += [torch.sin(v * 2.0 ** k), torch.cos(v * 2.0 ** k)]
        return torch.cat(parts, -1)

    off = 0
    mats = []
    for (o, i) in nets.shapes:
        w = nets.flat_ref[off:off + o * i].view(o, i)
        b = nets.flat_ref[off + o * i:off + o * i + o]
        mats.append((w, b))
        off += o * i + o
    xe = h16(embed(x, nfx))
    h = xe
    for i in range(D):
        w, b = mats[i]
        h = h16(torch.relu(h @ h16(w).t() + b))
        if i == skip:
            h = torch.cat([h, xe], -1)
    w, b = mats[D]
    sigma = h @ w[:1].t() + b[:1]                       # density row: fp32 weights on the CUDA cores
    feat = h16(h @ h16(w[1:]).t() + b[1:])
    ve = h16(embed(dirs / dirs.norm(dim=-1, keepdim=True), nfv))
    w, b = mats[D + 1]
    hr = h16(torch.relu(torch.cat([ve, feat], -1) @ h16(w).t() + b))
    w, b = mats[D + 2]
    return sigma, torch.sigmoid(hr @ h16(w).t() + b)


@pytest.mark.parametrize('name', list(nc.CASES))
def test_forward_matches_fp16_emulation(name):
    nets = build(name)
    x, dirs, _, _ = (torch.from_numpy(a).cuda() for a in nc.inputs(name))
    with torch.no_grad():
        sigma, rgb = nets(x, dirs)
        nets.flat_ref = nets.flat.detach()
        s2, c2 = _emulated(nets, x, dirs)
    assert float((sigma - s2).abs().max()) < 5e-4 and float((rgb - c2).abs().max()) < 2e-4


@pytest.mark.parametrize('name', list(nc.CASES))
def test_gradients_match_backprop_on_the_saved_activations(name):
    """manual backpropagation in torch (fp32) through the activations the forward kernel saved: same ReLU gates, so the
    data-gradient chain and the batch-contraction weight gradients must agree to rounding (fp16 gradient tiles: 5e-3)"""
    nets = build(name)
    nets.keep_save = True
    x, dirs, w_s, w_c = (torch.from_numpy(a).cuda() for a in nc.inputs(name))
    sigma, rgb = nets(x, dirs)
    ((sigma * w_s).sum() + (rgb * w_c).sum()).backward()
    torch.cuda.synchronize()
    save, n_cap = nets.last_save
    n = x.shape[0]
    d_x, nfx, nfv, W, D, skip, W_rad = nets.arch
    cols = nets.saved_columns()
    S = {k: nets.untile(save[c * n_cap:c * n_cap + n_cap * w], n_cap, w)[:n].float() for k, (c, w) in cols.items()}
    mats, off = [], 0
    for (o, i) in nets.shapes:
        mats.append(nets.flat.detach()[off:off + o * i].view(o, i))
        off += o * i + o
    h16 = lambda t: t.half().float()
    ev, ex = nets.embed_v, nets.embed_x
    exp = {}
    g = w_c * rgb.detach() * (1 - rgb.detach())
    exp[D + 2] = (g.t() @ S['hr'], g.sum(0))
    dz = (g @ h16(mats[D + 2])) * (S['hr'] > 0)
    exp[D + 1] = (torch.cat([dz.t() @ S['ve'][:, :ev], dz.t() @ S['feat']], 1), dz.sum(0))
    dfeat = dz @ h16(mats[D + 1][:, ev:])
    dsig = w_s[:, 0]
    exp[D] = (torch.cat([(dsig[None] @ S['h%d' % (D - 1)]), dfeat.t() @ S['h%d' % (D - 1)]], 0), torch.cat([dsig.sum()[None], dfeat.sum(0)]))
    dh = dfeat @ h16(mats[D][1:]) + dsig[:, None] * mats[D][:1]
    for i in range(D - 1, -1, -1):
        dz = dh * (S['h%d' % i] > 0)
        src = S['xe'][:, :ex] if i == 0 else (torch.cat([S['h%d' % (i - 1)], S['xe'][:, :ex]], 1) if i - 1 == skip else S['h%d' % (i - 1)])
        exp[i] = (dz.t() @ src, dz.sum(0))
        if i > 0:
            dh = dz @ h16(mats[i][:, :W])
    gflat, off = nets.flat.grad, 0
    for li, (o, i) in enumerate(nets.shapes):
        gw, gb = gflat[off:off + o * i].view(o, i), gflat[off + o * i:off + o * i + o]
        off += o * i + o
        ew, eb = exp[li]
        assert float((gw - ew).abs().max()) < 5e-3 * float(ew.abs().max()), ('weight', li)
        assert float((gb - eb).abs().max()) < 5e-3 * float(eb.abs().max()), ('bias', li)


def test_state_dict_round_trip_and_repack_after_update():
    nets = build('nerf')
    geo, rad = nets.reference_state()
    gs, rs = nc.weights('nerf')
    assert all(np.array_equal(geo[k].cpu().numpy(), gs[k]) for k in gs) and all(np.array_equal(rad[k].cpu().numpy(), rs[k]) for k in rs)
    x, dirs, _, _ = (torch.from_numpy(a).cuda() for a in nc.inputs('nerf'))
    with torch.no_grad():
        s0, _ = nets(x, dirs)
        nets.flat.mul_(0.5)                       # an optimiser step changes the parameters in place: blobs are re-packed
        s1, _ = nets(x, dirs)
    assert float((s0 - s1).abs().max()) > 1e-3


def test_unsupported_architecture_is_an_error():
    from arcnerf_b200.plugin.nerf_mlp import FusedNeRFNets
    with pytest.raises(RuntimeError):
        FusedNeRFNets(3, 10, 4, 128, 8, [4], 128)
