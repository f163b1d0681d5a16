"""GPU: the fused NGP pipeline (arcnerf_b200/ngp.py, libarcb200 kernels only) against the composed CPU oracle
(oracle/ngp.py) on the same rays, occupancy and random-init weights.
  bit-exact : per-ray sample counts, packed sample distances
  tolerance : rgb/depth/mask abs 2e-3 (fp16 activations in two MLPs), loss rel 2e-3,
              gradients rel 2e-2 of the largest entry; parameters after Adam steps abs 1e-3*lr-scale."""
import numpy as np
import pytest
import torch

from arcnerf_b200 import synth
from arcnerf_b200.ngp import NGPPipeline, NGP_DEFAULTS, init_params
from oracle.ngp import NGPOracle

pytestmark = pytest.mark.gpu


def setup(n_rays, seed, occupancy='lego', grid_scale=0.3, **over):
    cfg = dict(NGP_DEFAULTS); cfg.update(over)
    params = init_params(cfg, seed=seed, grid_scale=grid_scale)     # "trained-like" table so that sigma varies
    pipe = NGPPipeline(max_rays=n_rays, capacity=1 << 18, params=params, **over)
    orc = NGPOracle({k: v.numpy() for k, v in params.items()}, cfg)
    vol = synth.occupancy_volume(occupancy, n=cfg['n_grid'], seed=3)
    pipe.set_occupancy(vol); orc.set_occupancy(vol)
    o, d = synth.train_ray_batch(n_rays, seed=seed + 1)
    rng = np.random.RandomState(seed + 2)
    tgt = rng.uniform(0, 1, (n_rays, 3)).astype(np.float32)
    bkg = rng.uniform(0, 1, (n_rays, 3)).astype(np.float32)
    return pipe, orc, o, d, tgt, bkg


def dev(a):
    return torch.from_numpy(a).cuda()


def test_forward_matches_oracle():
    pipe, orc, o, d, tgt, bkg = setup(700, 0)
    rgb, depth, acc = pipe.forward(dev(o), dev(d), dev(bkg))
    res = orc.forward(o, d, bkg)
    counts = pipe.counts[:700].cpu().numpy()
    assert np.array_equal(counts, res['counts'])                              # bit-exact sample counts
    total = int(pipe.meta[0]); assert total == counts.sum() and int(pipe.meta[1]) == counts.max()
    assert np.array_equal(pipe.t[:total].cpu().numpy(), res['z'][res['m']])   # bit-exact sample distances
    np.testing.assert_allclose(rgb.cpu().numpy(), res['rgb'].numpy(), atol=2e-3, rtol=0)
    np.testing.assert_allclose(acc.cpu().numpy(), res['mask'].numpy(), atol=2e-3, rtol=0)
    np.testing.assert_allclose(depth.cpu().numpy(), res['depth'].numpy(), atol=1e-2, rtol=0)
    empty = counts == 0
    assert empty.any()
    np.testing.assert_array_equal(rgb.cpu().numpy()[empty], bkg[empty])


def test_training_step_gradients_and_adam():
    pipe, orc, o, d, tgt, bkg = setup(600, 5, lr=1e-2)
    loss = float(pipe.train_step(dev(o), dev(d), dev(tgt), dev(bkg), optimize=False))
    lo, grads, res = orc.loss_and_grads(o, d, tgt, bkg)
    assert abs(float(loss) - lo) <= 2e-3 * abs(lo)
    for name, g in (('rad', pipe.g_rad), ('geo', pipe.g_geo), ('table', pipe.g_table)):
        a, b = g.cpu().numpy(), grads[name]
        scale = np.abs(b).max()
        assert scale > 0
        err = np.abs(a - b)
        assert err.max() <= 2e-2 * scale, (name, err.max(), scale)
    # the table gradient touches exactly the entries the oracle touches
    assert np.array_equal(pipe.g_table.cpu().numpy() != 0, grads['table'] != 0) or \
        (np.logical_xor(pipe.g_table.cpu().numpy() != 0, grads['table'] != 0).mean() < 1e-4)
    # Adam: apply the step on both sides, several steps, compare parameters
    pipe.optimizer_step()
    assert float(pipe.g_table.abs().max()) == 0.0 and float(pipe.g_rad.abs().max()) == 0.0
    orc2 = NGPOracle(dict(table=orc.table, geo=orc.geo_p, rad=orc.rad_p), orc.cfg); orc2.set_occupancy(orc.volume)
    orc2.train_step(o, d, tgt, bkg)
    for a, b in ((pipe.rad_p, orc2.rad_p), (pipe.geo_p, orc2.geo_p), (pipe.table, orc2.table)):
        # the first Adam step moves every touched parameter by ~lr * sign(g): entries whose gradient is at the noise
        # floor may take the other sign (2*lr apart); everything else must agree
        err = np.abs(a.cpu().numpy() - b)
        assert (err > 2.5e-3).mean() < 5e-3 and err.max() <= 2.0 * 1e-2 + 1e-4
    assert torch.equal(pipe.table_h, pipe.table.half())
    for _ in range(3):
        l1 = float(pipe.train_step(dev(o), dev(d), dev(tgt), dev(bkg)))
    assert np.isfinite(l1) and l1 < float(loss)          # the loss goes down on a repeated batch


def test_render_chunks_and_capacity_clamp():
    pipe, orc, o, d, tgt, bkg = setup(512, 9)
    o2, d2 = synth.camera_rays(64, 48, synth.sphere_poses(1, seed=4)[0])
    b1 = torch.ones((1, 3), device='cuda')
    rgb, depth, acc = pipe.render(dev(o2), dev(d2), b1, chunk=512)
    assert rgb.shape == (64 * 48, 3) and torch.isfinite(rgb).all()
    # chunked == the same chunks run by hand with the same generator states
    pipe.rng = type(pipe.rng)()
    r2 = [pipe.forward(dev(o2[s:s + 512]), dev(d2[s:s + 512]), b1)[0].clone() for s in range(0, 64 * 48, 512)]
    pipe.rng = type(pipe.rng)()
    r3, _, _ = pipe.render(dev(o2), dev(d2), b1, chunk=512)
    assert torch.equal(torch.cat(r2), r3)
    # a capacity smaller than the sample count drops the overflow without touching memory out of bounds
    small = NGPPipeline(max_rays=512, capacity=4096, params=dict(table=pipe.table.cpu(), geo=pipe.geo_p.cpu(), rad=pipe.rad_p.cpu()))
    small.set_occupancy(synth.occupancy_volume('full'))
    rgb_s, _, _ = small.forward(dev(o[:512]), dev(d[:512]), None)
    assert int(small.meta[0]) > 4096 and torch.isfinite(rgb_s).all()
    l = small.train_step(dev(o[:512]), dev(d[:512]), dev(tgt[:512]))
    assert torch.isfinite(l).all()


def test_graph_replay_and_sampler_prefetch_equal_eager_steps():
    """Same batches through (a) plain eager steps, (b) eager steps with the next batch sampled on the side stream,
    (c) static slots + CUDA-graph replay: same sample counts, losses and parameters (atomics order aside)."""
    n = 512
    cfg = dict(NGP_DEFAULTS); cfg['lr'] = 1e-2
    params = init_params(cfg, seed=21, grid_scale=0.3)
    vol = synth.occupancy_volume('lego')
    batches = []
    for i in range(4):
        o, d = synth.train_ray_batch(n, seed=100 + i)
        rng = np.random.RandomState(200 + i)
        batches.append(tuple(torch.from_numpy(a).cuda() for a in
                             (o, d, rng.uniform(0, 1, (n, 3)).astype(np.float32), rng.uniform(0, 1, (n, 3)).astype(np.float32))))
    out = []
    for mode in ('eager', 'prefetch', 'graph'):
        pipe = NGPPipeline(max_rays=n, capacity=1 << 16, params={k: v.clone() for k, v in params.items()}, lr=1e-2)
        pipe.set_occupancy(vol)
        if mode == 'graph':
            pipe.enable_graphs(n)
        losses, counts = [], []
        for i, b in enumerate(batches):
            nxt = batches[i + 1] if i + 1 < len(batches) else None
            if mode == 'eager':
                l = pipe.train_step(*b)
            elif mode == 'prefetch':
                l = pipe.train_step(*b, next_rays=None if nxt is None else nxt[:2])
            else:
                l = pipe.train_step_graphed(b, next_batch=nxt)
            losses.append(float(l)); counts.append(int(pipe.meta[0]))
        torch.cuda.synchronize()
        out.append((losses, counts, pipe.table.clone(), pipe.geo_p.clone(), pipe.rad_p.clone(), int(pipe.step_dev), pipe.step_count))
    ref = out[0]
    assert ref[5] == 4 and ref[6] == 4
    for other in out[1:]:
        assert other[1] == ref[1]                                  # identical sample counts (same jitter stream)
        assert other[5] == 4 and other[6] == 4
        np.testing.assert_allclose(other[0], ref[0], rtol=2e-3)
        for a, b in zip(other[2:5], ref[2:5]):
            err = (a - b).abs()
            assert float((err > 2e-3).float().mean()) < 5e-3        # Adam sign flips at the noise floor only


def test_occupancy_refresh_through_volume_bound():
    """FgModel.optimize path: the pipeline's density query feeds VolumeBound.optimize, the refreshed bool volume is what the
    sampler marches through afterwards (fg_model.py:400-410, volume_bound.py:160-212)."""
    from arcnerf_b200 import synth
    from arcnerf_b200.ngp import NGP_DEFAULTS, NGPPipeline, init_params
    params = init_params(dict(NGP_DEFAULTS), seed=5, grid_scale=0.5)
    pipe = NGPPipeline(max_rays=512, capacity=1 << 16, params=params)
    pts = (torch.rand((3000, 3), device='cuda') - 0.5) * 1.9
    sig = pipe.forward_pts(pts, chunk=1024)
    sig_all = pipe.forward_pts(pts)
    assert sig.shape == (3000,) and torch.equal(sig, sig_all) and bool((sig > 0).all())
    assert torch.allclose(pipe.get_est_opacity(0.01, pts), sig * 0.01)
    assert bool(pipe.volume.all())
    pipe.update_occupancy(5)                       # not a refresh step
    assert bool(pipe.volume.all())
    pipe.update_occupancy(16)                      # warm-up refresh over all 128^3 voxels
    occ = float(pipe.volume.float().mean())
    vol = pipe.bound.volume
    dens = pipe.forward_pts(vol.get_volume_pts()).view(128, 128, 128)
    dt = vol.get_diag_len() / pipe.cfg['n_sample']
    thres = min(float(vol.get_voxel_opafield().clamp(min=0).mean()), 0.01)
    assert torch.equal(pipe.volume, vol.get_voxel_opafield() >= thres)
    # jittered sample vs voxel centre: the two occupancy estimates agree on most voxels
    agree = float(((dens * dt >= thres) == pipe.volume).float().mean())
    assert agree > 0.55 and 0.0 < occ <= 1.0      # random-init field: sub-voxel noise, but clearly better than chance
    o, d = synth.train_ray_batch(512, seed=9)
    t = torch.rand((512, 3), device='cuda')
    loss = pipe.train_step(torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda(), t)
    assert torch.isfinite(loss).all()


def test_bitfield_occupancy_gives_the_same_samples_and_step():
    """occupancy='bitfield' (sparse_volume_sampling_bit on the Morton bit field) and the bool volume describe the same
    scene: identical packed samples, identical forward, identical training step; the BitfieldBound refresh runs."""
    from arcnerf_b200 import synth
    from arcnerf_b200.ngp import NGP_DEFAULTS, NGPPipeline, init_params
    params = init_params(dict(NGP_DEFAULTS), seed=5, grid_scale=0.5)
    occ = synth.occupancy_volume('lego')
    pipes = []
    for kind in ('volume', 'bitfield'):
        p = NGPPipeline(max_rays=1024, capacity=1 << 17, params={k: v.clone() for k, v in params.items()}, occupancy=kind)
        p.set_occupancy(occ)
        pipes.append(p)
    o, d = synth.train_ray_batch(1024, seed=4)
    o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    t = torch.rand((1024, 3), device='cuda')
    outs = [p.forward(o, d, None) for p in pipes]
    a, b = pipes
    n = int(a.meta[0])
    assert n == int(b.meta[0]) and n > 0 and torch.equal(a.counts, b.counts)
    assert torch.equal(a.t[:n], b.t[:n]) and torch.equal(a.ray_id[:n], b.ray_id[:n]) and torch.equal(a.unit[:n], b.unit[:n])
    for x, y in zip(outs[0], outs[1]):
        assert torch.equal(x, y)
    la, lb = float(a.train_step(o, d, t)), float(b.train_step(o, d, t))
    assert abs(la - lb) <= 1e-5 * abs(la)
    b.update_occupancy(16)                                   # warm-up refresh through BitfieldBound
    count, ratio = b.bound.get_bitfield_count()
    assert 0.0 < ratio <= 1.0 and torch.equal(b.bits, b.bound.density_bitfield)
    assert torch.isfinite(b.train_step(o, d, t)).all()


def test_fused_field_forward_equals_separate_launches():
    """arcb200_field_fwd (encode + both heads in one launch) against arcb200_grid_fwd + arcb200_mlp_fwd_sh + arcb200_mlp_fwd_ex:
    same arithmetic, so every output and every tensor saved for the backward is identical"""
    pipe, _, o, d, tgt, bkg = setup(900, 11)
    assert pipe.fused_fwd
    o_d, d_d = dev(o), dev(d)
    out = {}
    for fused in (True, False):
        pipe.fused_fwd = fused
        pipe.rng = type(pipe.rng)()                 # same start jitter for both passes
        for ss in pipe.sets:
            ss.key = None
        for buf in (pipe.enc, pipe.rad_in, pipe.sigma, pipe.rgb_s, pipe.geo_act, pipe.rad_act):
            buf.zero_()
        pipe.forward(o_d, d_d, dev(bkg))
        torch.cuda.synchronize()
        n = int(pipe.meta[0])
        tiles = (n + 127) // 128
        out[fused] = dict(enc=pipe.enc[:n].clone(), rad_in=pipe.rad_in[:n].clone(), sigma=pipe.sigma[:n].clone(),
                          rgb_s=pipe.rgb_s[:n].clone(), rgb=pipe.rgb[:900].clone(),
                          # saved activations: rows of the last tile beyond n differ by design (the fused kernel encodes padding rows)
                          geo_act=pipe.geo_act[:(tiles - 1) * 128 * 64 * 2].clone(),
                          rad_act=pipe.rad_act[:(tiles - 1) * 2 * 128 * 64 * 2].clone())
    assert n > 128 * 20
    for k in out[True]:
        assert torch.equal(out[True][k], out[False][k]), k


def test_fused_field_forward_inference_writes_no_training_tensors():
    pipe, _, o, d, tgt, bkg = setup(300, 12)
    pipe.enc.fill_(7.0); pipe.rad_in.fill_(7.0)
    pipe.render(dev(o), dev(d), dev(bkg)[:1])
    torch.cuda.synchronize()
    assert float(pipe.enc.min()) == 7.0 and float(pipe.rad_in.min()) == 7.0
