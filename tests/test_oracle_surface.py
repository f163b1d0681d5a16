"""CPU: the numpy restatement of sphere_tracing / secant_root_finding (oracle/surface.py) against the outputs of the reference's own
functions (tests/golden/surface_golden.npz, made by tests/golden/make_golden_surface.py)."""
import os

import numpy as np
import pytest

import surface_cases as sc
from oracle import surface as osf

G = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'surface_golden.npz'))


@pytest.mark.parametrize('name', list(sc.CASES))
def test_restatement_matches_reference_functions(name):
    method, fname, kw = sc.CASES[name]
    o, d, near, far = sc.rays()
    kw = dict(kw)
    for k in ('near', 'far'):
        if kw[k] == 'per_ray':
            kw[k] = near if k == 'near' else far
    z, pts, mask = getattr(osf, method)(o, d, sc.field(fname, np), **kw)
    assert np.array_equal(mask, G[name + '/mask'])
    # the analytic field itself differs in the last bit between numpy and torch (torch's vectorised CPU kernels contract x*x + y*y
    # into a fused multiply-add), and a traced distance is a sum of up to n_iter field values: a few ulps at |z| ~ 6
    assert np.allclose(z, G[name + '/zvals'], rtol=0, atol=1e-5)
    assert np.allclose(pts, G[name + '/pts'], rtol=0, atol=2e-5)
    assert 20 < mask.sum() < len(mask)
