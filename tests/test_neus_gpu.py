"""NeuS-NGP foreground model (plugin/neus_model.py; BASELINE.json configs[2]) on the GPU against the composed CPU oracle
(oracle/neus.py: every piece pinned to the reference, see its header), and a training run with the eikonal term, i.e. with the
second-order backward through the hash grid and the SDF network.
Tolerances: sample counts / distances bit-exact (same sampler as the NGP tests); rgb / mask abs 5e-3, depth 2e-2, normal map
5e-2 (fp16 hash-grid features enter a 100-beta softplus network whose input gradient is taken)."""
import numpy as np
import pytest
import torch

from arcnerf_b200 import synth

pytestmark = pytest.mark.gpu

CFG = dict(
    obj_bound=dict(volume=dict(n_grid=128, side=2.0), epoch_optim=16, epoch_optim_warmup=256, ray_sample_acc=True, ray_sample_fix_step=True,
                   near_distance=0.2, opa_thres=0.01, bkg_color=[1.0, 1.0, 1.0]),
    rays=dict(radius_bound=1.5, n_sample=1024, n_iter=4, inverse_linear=False, perturb=True, add_inf_z=False, noise_std=0.0, white_bkg=True),
    params=dict(speed_factor=10, init_var=0.05, anneal_end=50000), chunk_rays=32768, chunk_pts=1048576,
    geometry=dict(W=64, D=1, use_bias=False, skips=[], W_feat=16, geometric_init=True, radius_init=0.5, weight_norm=True,
                  skip_reduce_output=True, norm_skip=True, act_cfg=dict(type='softplus', beta=100),
                  encoder=dict(type='HashGridEmbedder', input_dim=3, n_freqs=0, side=2.0, n_levels=16, n_feat_per_entry=2, hashmap_size=19,
                               base_res=16, max_res=2048, include_input=False, backend='tcnn', dtype='torch.float32')),
    radiance=dict(type='FusedMLPRadianceNet', mode='pvnf', W=64, D=2, W_feat_in=16, weight_norm=True,
                  encoder=dict(pts=dict(type='FreqEmbedder', input_dim=3, n_freqs=0),
                               view=dict(type='SHEmbedder', input_dim=3, n_freqs=4, include_input=False, backend='tcnn'))))


def build(seed=0, table_scale=0.05):
    from arcnerf_b200.plugin import volume_func
    from arcnerf_b200.plugin.neus_model import Neus
    volume_func.reset_rng()                   # the module-level sampler generator (common.h:21-23) as after a fresh import
    torch.manual_seed(seed)
    model = Neus(CFG)
    with torch.no_grad():                     # a non-trivial hash table so that the SDF varies in space
        g = torch.Generator().manual_seed(seed + 1)
        p = model.geo_net.embed_fn.hashgrid_encode_tcnn.params
        p.copy_(((torch.rand(p.shape, generator=g) * 2 - 1) * table_scale).to(p.device))
    vol = synth.occupancy_volume('lego')
    model.obj_bound.volume.bitfield.copy_(torch.from_numpy(vol).to(model.inv_s.device))
    return model, vol


def test_forward_matches_composed_oracle():
    from oracle import fields, neus as on
    from oracle.sampler import Sampler
    model, vol = build()
    n = 160
    o, d = synth.train_ray_batch(n, seed=4)
    feed = {'rays_o': torch.from_numpy(o).cuda(), 'rays_d': torch.from_numpy(d).cuda()}
    with torch.no_grad():
        out = model(feed, inference_only=True)
    # ---- oracle: same sampler state (first launch of the volume module's generator), same weights
    smp = Sampler('oracle')
    aabb = synth.aabb_of_side(2.0)
    near, far, _, hit = smp.aabb_intersection(o, d, aabb[None])
    dt = np.float32(2.0 * np.sqrt(3.0)) / np.float32(1024)
    z, m = smp.sparse_volume_sampling(o, d, near, far, 1024, dt, aabb, 128, vol, 0.2, smp.rng())
    valid = hit[:, 0] & m.any(1)
    pmax = max(2, int(m.sum(1).max()))
    z, m = z[:, :pmax], m[:, :pmax]
    geo = model.geo_net
    w0 = on.weight_norm(geo.layers[0].weight_g.detach().cpu().numpy().astype(np.float64), geo.layers[0].weight_v.detach().cpu().numpy().astype(np.float64))
    w1 = on.weight_norm(geo.layers[1].weight_g.detach().cpu().numpy().astype(np.float64), geo.layers[1].weight_v.detach().cpu().numpy().astype(np.float64))
    table = geo.embed_fn.hashgrid_encode_tcnn.params.detach().cpu().numpy()
    pls = float(geo.embed_fn.per_level_scale)
    geom = fields.grid_geometry(16, 19, 16, pls)
    rad_p = model.radiance_net.layers.params.detach().cpu().numpy()
    ref = on.forward(o[valid], d[valid], z[valid], m[valid], table, geom, aabb, w0, w1, rad_p, float(model.inv_s.detach()), n_sample=1024)
    assert ref['n_valid'] > 3000 and valid.sum() > 40
    got = {k: out[k].cpu().numpy() for k in ('rgb', 'depth', 'mask', 'normal')}
    assert np.abs(got['rgb'][valid] - ref['rgb']).max() < 5e-3
    assert np.abs(got['mask'][valid] - ref['mask']).max() < 5e-3
    assert np.abs(got['depth'][valid] - ref['depth']).max() < 2e-2
    assert np.abs(got['normal'][valid] - ref['normal']).max() < 5e-2
    # rays that miss the occupancy: background colour, far depth, empty mask, default normal (fg_model.py:320-387)
    assert np.allclose(got['rgb'][~valid], 1.0) and np.allclose(got['depth'][~valid], 10.0) and np.allclose(got['mask'][~valid], 0.0)
    assert np.allclose(got['normal'][~valid], np.array([0.0, 1.0, 0.0]))


def test_training_with_eikonal_term_and_occupancy_refresh():
    from arcnerf_b200.plugin.neus_model import eikonal_loss
    model, _ = build(seed=2, table_scale=1e-4)
    n = 512
    o, d = synth.train_ray_batch(n, seed=9)
    feed = {'rays_o': torch.from_numpy(o).cuda(), 'rays_d': torch.from_numpy(d).cuda(), 'bkg_color': torch.rand((n, 3)).cuda()}
    target = torch.rand((n, 3), generator=torch.Generator().manual_seed(3)).cuda() * 0.6 + 0.2
    opt = torch.optim.Adam(model.parameters(), lr=2e-3)
    losses, eiks = [], []
    for it in range(10):
        out = model(feed, cur_epoch=it)
        assert set(out) >= {'rgb', 'depth', 'mask', 'normal', 'normal_pts', 'params'}
        eik = eikonal_loss(out['normal_pts'])
        loss = torch.nn.functional.huber_loss(out['rgb'], target, delta=0.1) * 5.0 + 0.1 * eik
        opt.zero_grad()
        loss.backward()
        for name, p in model.named_parameters():
            assert p.grad is None or bool(torch.isfinite(p.grad).all()), name
        opt.step()
        losses.append(float(loss.detach())); eiks.append(float(eik.detach()))
    assert model.geo_net.layers[0].weight_v.grad is not None and float(model.geo_net.layers[0].weight_v.grad.abs().sum()) > 0
    assert model.inv_s.grad is not None
    assert losses[-1] < losses[0], losses
    # the SDF's occupancy refresh (warm-up sweep over every voxel) runs and keeps a sane volume
    model.optimize(cur_epoch=16)
    occ = float(model.obj_bound.volume.bitfield.float().mean())
    assert 0.0 <= occ <= 1.0


def test_vanilla_neus_upsampling_and_render_match_reference_model():
    """The frequency-embedded, sphere-bounded NeuS of configs/models/neus.yaml (narrower networks, tests/neus_cases.py) with the
    reference model's own weights: the hierarchical up-sampling (upsample_zvals, neus_model.py:106-171) and the rendered outputs
    against the reference model run on CPU (tests/golden/make_golden_neus.py)."""
    import os
    import neus_cases as nc
    from arcnerf_b200.plugin.neus_model import Neus
    G = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'neus_golden.npz'))
    model = Neus(nc.CFG)
    sd = {k[3:]: torch.from_numpy(G[k]) for k in G.files if k.startswith('sd/')}
    missing, unexpected = model.load_state_dict(sd, strict=False)
    assert not unexpected and all('embed' in k or 'freq' in k for k in missing), (missing, unexpected)
    o, d = nc.rays()
    ro, rd = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    with torch.no_grad():
        near, far, mask_rays = model.obj_bound.get_near_far_from_rays({'rays_o': ro, 'rays_d': rd})
        zvals, mask_pts = model.obj_bound.get_zvals_from_near_far(near, far, 32, True, False, False, rays_o=ro, rays_d=rd)
        assert mask_pts is None and np.array_equal(mask_rays.cpu().numpy(), G['mask_rays'])
        assert np.allclose(zvals.cpu().numpy(), G['zvals'], atol=1e-5)
        z_up, m_up = model.upsample_zvals(ro, rd, zvals, None, inference_only=True)
        out = model({'rays_o': ro, 'rays_d': rd}, inference_only=True)
    assert z_up.shape == (nc.N_RAYS, 64) and m_up is None
    hit = G['mask_rays']
    # the SDF network runs in fp32 on both sides; an inverse-CDF sample moves with the CDF, which is steep near the surface
    dz = np.abs(z_up.cpu().numpy() - G['z_up'])[hit]
    assert np.median(dz) < 1e-5 and dz.max() < 5e-3
    for k, tol in (('rgb', 5e-3), ('mask', 5e-3), ('depth', 2e-2), ('normal', 3e-2)):
        assert np.abs(out[k].cpu().numpy() - G['out/' + k])[hit].max() < tol, k
