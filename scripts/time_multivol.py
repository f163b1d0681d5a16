"""Times the cascaded (multi-volume) march, arcb200_sparse_sampling_in_multivol_bitfield, on 4096 unbounded-capture rays, 1024
samples per ray, 4 cascades of a 128^3 Morton bit field at 10 % random occupancy -> one JSON line (profiles/r02_multivol_march.json)."""
import json

import numpy as np
import torch

from arcnerf_b200.plugin import multivol_func as mv

torch.manual_seed(0)
n, n_pts, n_grid, n_cascade = 4096, 1024, 128, 4
rs = np.random.RandomState(1)
o = rs.normal(size=(n, 3)); o = o / np.linalg.norm(o, axis=1, keepdims=True) * 0.8
d = rs.normal(size=(n, 3)); d = d / np.linalg.norm(d, axis=1, keepdims=True)
o, d = torch.from_numpy(o.astype(np.float32)).cuda(), torch.from_numpy(d.astype(np.float32)).cuda()
inner = torch.tensor([[-1.0, 1.0]] * 3, device='cuda')
outer = inner * 2 ** (n_cascade - 1)
bits = (torch.rand(n_grid ** 3 // 8 * (n_cascade - 1), device='cuda') < 0.1).to(torch.uint8) * 0x11
near = torch.full((n, 1), 0.05, device='cuda')
far = torch.full((n, 1), 12.0, device='cuda')
args = (o, d, near, far, n_pts, 1.0 / 256, 2 * np.sqrt(3) / 1024, 2 * np.sqrt(3) * 8 / 1024, inner, outer, n_grid, n_cascade, bits, 0.05, False)
for _ in range(3):
    z, m = mv.sparse_sampling_in_multivol_bitfield(*args)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
for _ in range(20):
    z, m = mv.sparse_sampling_in_multivol_bitfield(*args)
e1.record()
torch.cuda.synchronize()
print(json.dumps(dict(kernel='march_multivol_kernel', rays=n, n_pts=n_pts, cascades=n_cascade, samples=int(m.sum()),
                      ms_per_call=e0.elapsed_time(e1) / 20, output_bytes=n * n_pts * 5)))
