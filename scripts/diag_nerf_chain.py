import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import nerf_full_cases as nc
from test_nerf_chain_gpu import build, _emulated
for name in nc.CASES:
    nets = build(name)
    x, dirs, w_s, w_c = (torch.from_numpy(a).cuda() for a in nc.inputs(name))
    sigma, rgb = nets(x, dirs)
    ((sigma * w_s).sum() + (rgb * w_c).sum()).backward()
    nets.flat_ref = nets.flat.detach().clone().requires_grad_(True)
    s2, c2 = _emulated(nets, x, dirs)
    ((s2 * w_s).sum() + (c2 * w_c).sum()).backward()
    g, gr = nets.flat.grad, nets.flat_ref.grad
    print(name, 'fwd err sigma %.2e rgb %.2e' % (float((sigma - s2).abs().max()), float((rgb - c2).abs().max())))
    off = 0
    for li, (o, i) in enumerate(nets.shapes):
        a, b = g[off:off + o * i].view(o, i), gr[off:off + o * i].view(o, i)
        ab, bb = g[off + o * i:off + o * i + o], gr[off + o * i:off + o * i + o]
        e = (a - b).abs()
        print('  layer %2d (%3d x %3d) W: max err/max %.2e  relL2 %.2e  worst col %d row %d | b: %.2e' % (
            li, o, i, float(e.max() / b.abs().max()), float((a - b).norm() / b.norm()), int(e.max(0).values.argmax()), int(e.max(1).values.argmax()),
            float((ab - bb).abs().max() / bb.abs().max())))
        off += o * i + o
