"""NGPFgModel (plugin/fg_model.py): the reference's FgModel.forward contract on the packed pipeline - output keys,
defaults for rays that miss the bound, chunked inference, autograd gradients equal to the fused step's gradients, and an
external torch optimiser driving the module."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _batch(n, seed):
    from arcnerf_b200 import synth
    o, d = synth.train_ray_batch(n, seed=seed)
    rng = np.random.RandomState(seed + 50)
    return (torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda(),
            torch.from_numpy(rng.uniform(0, 1, (n, 3)).astype(np.float32)).cuda(),
            torch.from_numpy(rng.uniform(0, 1, (n, 3)).astype(np.float32)).cuda())


def _model(max_rays=512, **kw):
    from arcnerf_b200 import synth
    from arcnerf_b200.ngp import NGP_DEFAULTS, init_params
    from arcnerf_b200.plugin.fg_model import NGPFgModel
    params = init_params(dict(NGP_DEFAULTS), seed=7, grid_scale=0.5)
    cfgs = dict(rays=dict(n_sample=1024), obj_bound=dict(volume=dict(n_grid=128, side=2.0), epoch_optim=16,
                                                       epoch_optim_warmup=256, near_distance=0.2),
                render=dict(bkg_color=[1.0, 1.0, 1.0], depth_far=10.0))
    m = NGPFgModel(cfgs, params={k: v.clone() for k, v in params.items()}, max_rays=max_rays, capacity=1 << 17, **kw)
    m.pipe.set_occupancy(synth.occupancy_volume('lego'))
    return m


def test_forward_contract_and_invalid_ray_defaults():
    m = _model()
    o, d, t, b = _batch(512, 1)
    o[:5] = torch.tensor([5.0, 5.0, 5.0], device='cuda')           # these rays look away from the box
    d[:5] = torch.tensor([0.0, 0.0, 1.0], device='cuda')
    out = m({'rays_o': o, 'rays_d': d, 'bkg_color': b})
    assert set(out) == {'rgb_coarse', 'depth_coarse', 'mask_coarse'}
    assert out['rgb_coarse'].shape == (512, 3) and out['depth_coarse'].shape == (512,) and out['rgb_coarse'].requires_grad
    assert torch.allclose(out['rgb_coarse'][:5], b[:5]) and bool((out['depth_coarse'][:5] == 10.0).all())
    assert bool((out['mask_coarse'][:5] == 0).all())
    inf = m({'rays_o': o, 'rays_d': d, 'bkg_color': None}, inference_only=True)
    assert set(inf) == {'rgb', 'depth', 'mask'} and not inf['rgb'].requires_grad
    assert torch.allclose(inf['rgb'][:5], torch.ones((5, 3), device='cuda'))          # configured white background
    # chunked inference == one chunk per call
    big_o, big_d = torch.cat([o, o.flip(0)]), torch.cat([d, d.flip(0)])
    with torch.no_grad():
        whole = m({'rays_o': big_o, 'rays_d': big_d, 'bkg_color': None}, inference_only=True)
    assert whole['rgb'].shape == (1024, 3)
    # every launch of the march draws a new start jitter (as the reference's kernels do), so chunks agree up to sampling noise
    assert float((whole['rgb'][:512] - inf['rgb']).abs().mean()) < 0.03
    assert float((whole['rgb'][512:] - inf['rgb'].flip(0)).abs().mean()) < 0.03
    assert torch.allclose(whole['rgb'][:5], inf['rgb'][:5]) and bool((whole['depth'][:5] == 10.0).all())
    with pytest.raises(RuntimeError, match='one chunk'):
        m({'rays_o': big_o, 'rays_d': big_d, 'bkg_color': None})


def test_autograd_gradients_equal_the_fused_step():
    from arcnerf_b200.plugin.ray_helper import sphere_ray_intersection  # noqa: F401  (library present)
    a, b_ = _model(), _model()
    o, d, t, bkg = _batch(512, 2)
    # module path: forward -> Huber loss in torch -> backward
    out = a({'rays_o': o, 'rays_d': d, 'bkg_color': bkg})
    loss = 3000.0 * torch.nn.functional.huber_loss(out['rgb_coarse'], t, delta=0.1) / 0.1
    loss.backward()
    # fused path, no optimiser step: same march (same generator state), same gradients
    lf = b_.pipe.train_step(o, d, t, bkg, optimize=False)
    assert abs(float(loss.detach()) - float(lf)) <= 2e-4 * abs(float(lf))
    for g_mod, g_fused in ((a.table.grad, b_.pipe.g_table), (a.geo_params.grad, b_.pipe.g_geo), (a.rad_params.grad, b_.pipe.g_rad)):
        scale = float(g_fused.abs().max())
        assert scale > 0 and float((g_mod - g_fused).abs().max()) <= 2e-3 * scale
    # depth / mask gradients flow as well
    a.zero_grad()
    out = a({'rays_o': o, 'rays_d': d, 'bkg_color': bkg})
    (out['depth_coarse'].sum() + out['mask_coarse'].sum()).backward()
    assert float(a.table.grad.abs().max()) > 0


def test_external_optimizer_drives_the_module():
    m = _model()
    opt = torch.optim.Adam(m.parameters(), lr=1e-2, eps=1e-15)
    o, d, t, bkg = _batch(512, 3)
    losses = []
    for _ in range(12):
        opt.zero_grad()
        out = m({'rays_o': o, 'rays_d': d, 'bkg_color': bkg})
        loss = torch.nn.functional.mse_loss(out['rgb_coarse'], t)
        loss.backward()
        opt.step()                      # writes the fp32 masters; the fp16 table / packed weights follow at the next forward
        losses.append(float(loss))
    assert losses[-1] < 0.92 * losses[0]        # random per-ray targets: only part of the loss is reducible
    # occupancy refresh and the fused step keep working on the same module
    m.optimize(16)
    l2 = m.train_step(o, d, t, bkg)
    assert torch.isfinite(l2).all() and 0.0 < float(m.pipe.volume.float().mean()) <= 1.0
