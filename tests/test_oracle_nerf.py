"""oracle/nerf.py (torch-CPU restatement of the NeRF foreground, the NeRF++ background and their blend) against outputs of the
REFERENCE's own FullModel run on CPU (tests/golden/models_golden.npz, tests/golden/make_golden_models.py)."""
import os

import numpy as np
import torch

import model_cases as mc
import nerf_full_cases as nc
from oracle import nerf as on

G = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'models_golden.npz'))


def _nets(name, seed, **kw):
    gs, rs = nc.weights(name, seed)
    return on.Nets(gs, rs, **kw)


def test_nerf_foreground_matches_reference_model():
    o, d = (torch.from_numpy(a) for a in mc.rays())
    c, f = _nets('nerf', mc.SEEDS['fg_coarse']), _nets('nerf', mc.SEEDS['fg_fine'])
    with torch.no_grad():
        out_c, out_f = on.nerf_forward(c, f, o, d, mc.NERF_CFG['rays'])
    for k in ('rgb', 'depth', 'mask'):
        assert np.allclose(out_c[k].numpy(), G['nerf/train/%s_coarse' % k], atol=2e-5), k
        assert np.allclose(out_f[k].numpy(), G['nerf/train/%s_fine' % k], atol=2e-5), k
        assert np.allclose(out_f[k].numpy(), G['nerf/inference/%s' % k], atol=2e-5), k


def test_nerfpp_full_model_matches_reference_model():
    o, d = (torch.from_numpy(a) for a in mc.rays())
    fc, ff = _nets('nerf', mc.SEEDS['fg_coarse']), _nets('nerf', mc.SEEDS['fg_fine'])
    bc, bf = _nets('nerfpp_bkg', mc.SEEDS['bkg_coarse']), _nets('nerfpp_bkg', mc.SEEDS['bkg_fine'])
    with torch.no_grad():
        fg_c, fg_f = on.nerf_forward(fc, ff, o, d, mc.NERFPP_FG_CFG['rays'])
        bk_c, bk_f = on.nerfpp_forward(bc, bf, o, d, mc.NERFPP_BKG_CFG['rays'])
    out = on.blend(fg_c, fg_f, bk_c, bk_f)
    for k in ('rgb', 'depth', 'mask'):
        for st in ('coarse', 'fine'):
            assert np.allclose(out['%s_%s' % (k, st)].numpy(), G['nerfpp/train/%s_%s' % (k, st)], atol=3e-5), (k, st)
        assert np.allclose(out[k + '_fine'].numpy(), G['nerfpp/inference/' + k], atol=3e-5), k
    assert float(np.abs(G['nerfpp/train/rgb_fine'] - fg_f['rgb'].numpy()).max()) > 1e-2        # the background does contribute
