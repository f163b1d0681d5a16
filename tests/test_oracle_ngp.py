"""CPU: the torch-CPU (threaded) restatement used for the timed baseline equals the numpy oracle, and the composed
NGP oracle step runs and produces consistent gradients."""
import numpy as np
import torch

from arcnerf_b200 import synth
from oracle import fields
from oracle.ngp import NGPOracle, TorchGrid

SMALL = dict(n_levels=6, n_features=2, log2_hashmap_size=12, base_res=8, max_res=128, geo_width=64, geo_hidden=1, w_feat=16,
             rad_width=64, rad_hidden=2, sh_degree=4, n_grid=32, side=2.0, n_sample=128, near_distance=0.2, lr=1e-2,
             beta1=0.9, beta2=0.999, eps=1e-15, weight_decay=1e-6, huber_delta=0.1, loss_weight=3000.0, trunc_exp_clip=15.0)


def small_params(cfg, seed=0):
    rng = np.random.RandomState(seed)
    pls = float(torch.exp(torch.log(torch.tensor(cfg['max_res'] / cfg['base_res'])) / (float(cfg['n_levels']) - 1)))
    geom = fields.grid_geometry(cfg['n_levels'], cfg['log2_hashmap_size'], cfg['base_res'], pls)
    n_table = int(geom[0][-1]) * cfg['n_features']

    def mlp(i, w, h, o):
        shapes, _, _ = fields.mlp_shapes(i, w, h, o)
        return np.concatenate([rng.uniform(-1, 1, a * b) * np.sqrt(6.0 / (a + b)) for a, b in shapes]).astype(np.float32)

    return dict(table=rng.uniform(-0.5, 0.5, n_table).astype(np.float32),
                geo=mlp(cfg['n_levels'] * cfg['n_features'], cfg['geo_width'], cfg['geo_hidden'], cfg['w_feat']),
                rad=mlp(cfg['w_feat'] + cfg['sh_degree'] ** 2, cfg['rad_width'], cfg['rad_hidden'], 3)), geom


def test_torch_grid_equals_numpy_grid():
    params, geom = small_params(SMALL)
    rng = np.random.RandomState(1)
    x = rng.uniform(0, 1, (500, 3)).astype(np.float32)
    tg = TorchGrid(geom)
    th = torch.from_numpy(params['table'].astype(np.float16).astype(np.float32).reshape(-1, 2))
    a = tg.encode(torch.from_numpy(x), th).numpy()
    b = fields.grid_encode(x, params['table'].reshape(-1, 2), geom, 2)
    assert np.abs(a - b).max() <= 2.0 ** -10 * 0.5 and (a == b).mean() > 0.99
    g = rng.normal(size=a.shape).astype(np.float32)
    ga = tg.backward(torch.from_numpy(x), torch.from_numpy(g), th.shape[0]).numpy()
    gb = fields.grid_backward_table(x, g, geom, 2)
    np.testing.assert_allclose(ga, gb, rtol=1e-4, atol=1e-5)


def test_ngp_oracle_step():
    params, _ = small_params(SMALL)
    orc = NGPOracle(params, SMALL)
    orc.set_occupancy(synth.occupancy_volume('lego', n=32, seed=3))
    o, d = synth.train_ray_batch(200, seed=2)
    rng = np.random.RandomState(3)
    tgt = rng.uniform(0, 1, (200, 3)).astype(np.float32)
    bkg = rng.uniform(0, 1, (200, 3)).astype(np.float32)
    loss, grads, res = orc.loss_and_grads(o, d, tgt, bkg)
    assert res['counts'].sum() > 500 and np.isfinite(loss)
    assert all(np.isfinite(g).all() and np.abs(g).max() > 0 for g in grads.values())
    # rays without samples show the background
    empty = res['counts'] == 0
    assert empty.any()
    np.testing.assert_array_equal(res['rgb'].detach().numpy()[empty], bkg[empty])
    # directional derivative of the loss along the radiance-head weights (fp16 weight rounding bounds accuracy)
    orc2 = NGPOracle(params, SMALL); orc2.set_occupancy(orc.volume)
    direction = rng.normal(size=params['rad'].shape).astype(np.float32)
    eps = 4e-3
    vals = []
    for s in (+1, -1):
        p2 = dict(params); p2['rad'] = params['rad'] + s * eps * direction
        o2 = NGPOracle(p2, SMALL); o2.set_occupancy(orc.volume)
        vals.append(o2.loss_and_grads(o, d, tgt, bkg)[0])
    fd = (vals[0] - vals[1]) / (2 * eps)
    an = float((grads['rad'] * direction).sum())
    assert abs(fd - an) <= 0.1 * abs(an) + 1.0
    l0 = orc2.train_step(o, d, tgt, bkg)
    assert np.isfinite(l0) and not np.array_equal(orc2.rad_p, params['rad'])


def test_oracle_config_and_init_match_the_product():
    """bench.py's CPU arm takes its configuration and initial parameters from oracle/ (it must not import the product); they have
    to be the same values and bytes the GPU arm uses."""
    import torch
    from arcnerf_b200.ngp import NGP_DEFAULTS, init_params as product_init
    from oracle.ngp import NGP_CONFIG, init_params as oracle_init
    assert NGP_CONFIG == NGP_DEFAULTS
    a, b = product_init(dict(NGP_DEFAULTS), seed=5), oracle_init(dict(NGP_CONFIG), seed=5)
    for k in ('table', 'geo', 'rad'):
        assert torch.equal(a[k], b[k]), k
