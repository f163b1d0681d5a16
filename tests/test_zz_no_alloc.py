"""GPU: arcb200_march_packed_static - the packed march for hosts that cannot supply scratch memory (a sampler inside
another engine's frame callback, where cudaMalloc would stall the engine's streams; a plug-in whose host owns all
device memory and hands it only the call's buffers).  Its scratch is a fixed device array inside the library.

Checked here, through the C ABI: two calls in a row - the second one smaller, so it runs over whatever the first left
in the shared array - give exactly the counts, offsets, totals, sample distances, positions and ray ids of
arcb200_march_packed with caller-provided scratch (bit-exact, the tolerance of the other sampler tests); a request
beyond the documented size limit is refused with a clear error.  This file sorts last so that a run stopping at the
first failure has already covered everything else; it also passes when run on its own in a fresh process.
"""
import numpy as np
import pytest
import torch

import cases
from arcnerf_b200 import _lib

pytestmark = pytest.mark.gpu


def T(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def run_both(case, n_rays, dev, seed_advance):
    c = cases.march_case(case)
    o, d = T(c['o'][:n_rays], dev), T(c['d'][:n_rays], dev)
    aabb = T(c['aabb'], dev)
    vol = T(c['vol'], dev)
    n = o.shape[0]
    near = torch.empty((n, 1), device=dev); far = torch.empty((n, 1), device=dev)
    hit = torch.empty((n, 1), dtype=torch.uint8, device=dev)
    st = _lib.current_stream()
    _lib.call('arcb200_aabb_intersection', o.data_ptr(), d.data_ptr(), aabb.data_ptr(), near.data_ptr(), far.data_ptr(),
              None, hit.data_ptr(), n, 1, st)
    rng = _lib.Rng()
    for _ in range(seed_advance):
        rng.advance()
    cap = n * c['n_pts']
    outs = []
    for static in (False, True):
        counts = torch.full((n,), -7, dtype=torch.int32, device=dev)
        offsets = torch.full((n + 1,), -7, dtype=torch.int32, device=dev)
        meta = torch.full((2,), -7, dtype=torch.int32, device=dev)
        t = torch.full((cap,), -1.0, device=dev)
        xyz = torch.full((cap, 3), -1.0, device=dev)
        unit = torch.full((cap, 3), -1.0, device=dev)
        rid = torch.full((cap,), -1, dtype=torch.int32, device=dev)
        common = (0, o.data_ptr(), d.data_ptr(), near.data_ptr(), far.data_ptr(), c['n_pts'], float(c['dt']),
                  aabb.data_ptr(), c['n_grid'], vol.data_ptr(), float(c['near_distance']), n, rng.state, rng.inc)
        tail = (counts.data_ptr(), offsets.data_ptr(), meta.data_ptr(), cap, t.data_ptr(), xyz.data_ptr(), rid.data_ptr(),
                unit.data_ptr(), st)
        if static:
            _lib.call('arcb200_march_packed_static', *common, *tail)
        else:
            nbytes = _lib.load().arcb200_march_packed_scratch_bytes(n, c['n_pts'])
            scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            _lib.call('arcb200_march_packed', *common, scratch.data_ptr(), *tail)
        torch.cuda.synchronize()
        outs.append((counts, offsets, meta, t, xyz, unit, rid))
    return outs


def test_static_scratch_march_matches_caller_scratch_march(cuda):
    # first call: the NGP training batch, which fills the static array to its documented limit (4096 x 1024 slots)
    assert _lib.load().arcb200_march_packed_static_max_slots() == 4096 * 1024
    ref, got = run_both('ngp4096', 4096, cuda, seed_advance=0)
    assert int(ref[2][0]) > 100000
    for a, b in zip(ref, got):
        assert torch.equal(a, b)
    # second call straight after it: fewer, different rays and another jitter - stale rows of the first call remain
    ref, got = run_both('small32', 200, cuda, seed_advance=3)
    assert int(ref[2][0]) > 0
    for a, b in zip(ref, got):
        assert torch.equal(a, b)
    # beyond the limit: refused, nothing launched
    c = cases.march_case('ngp4096')
    n = 4097
    z = torch.zeros((n, 3), device=cuda)
    i32 = torch.zeros(n + 1, dtype=torch.int32, device=cuda)
    with pytest.raises(RuntimeError, match='exceeds the static scratch'):
        _lib.call('arcb200_march_packed_static', 0, z.data_ptr(), z.data_ptr(), z.data_ptr(), z.data_ptr(), 1024, 0.01,
                  z.data_ptr(), 128, z.data_ptr(), 0.0, n, 1, 3, i32.data_ptr(), i32.data_ptr(), i32.data_ptr(), 16,
                  z.data_ptr(), None, None, None, _lib.current_stream())
