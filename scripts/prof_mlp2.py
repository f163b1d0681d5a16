"""Backward-kernel timing of the two NGP heads with saved activations, for n rows (tiles-per-CTA experiments)."""
import sys
import torch
sys.path.insert(0, '.')
from arcnerf_b200 import functional as F

dev = torch.device('cuda:0')
for n in [int(a) for a in sys.argv[1:]] or [147624]:
    for dims, acts in (((32, 64, 1, 16), ('relu', 'none')), ((32, 64, 2, 3), ('relu', 'sigmoid'))):
        spec = F.MlpSpec(*dims, *acts)
        params = (torch.rand(spec.n_params, device=dev) - 0.5) * 0.5
        blob = F.mlp_pack(params, spec)
        x = (torch.rand((n, 32), device=dev) - 0.5).half()
        dy = torch.randn((n, dims[3]), device=dev) * 1e-2
        out = torch.empty((n, dims[3]), dtype=torch.float32, device=dev)
        act = F.mlp_act_save_buffer(n, spec, dev)
        dW = torch.zeros(spec.n_params, device=dev)
        dx = torch.empty((n, 32), dtype=torch.float32, device=dev)
        for _ in range(3):
            F.mlp_forward(x, blob, spec, out=out, act_save=act)
            F.mlp_backward(x, blob, spec, dy, dW=dW, dx=dx, act_saved=act, y_saved=out)
        torch.cuda.synchronize()
        a, b, c = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        torch.cuda._sleep(4000000)      # let the CPU run ahead: the events then see GPU time only
        a.record()
        for _ in range(20):
            F.mlp_forward(x, blob, spec, out=out, act_save=act)
        b.record()
        for _ in range(20):
            F.mlp_backward(x, blob, spec, dy, dW=dW, dx=dx, act_saved=act, y_saved=out)
        c.record()
        torch.cuda.synchronize()
        print('n=%d tiles=%d' % (n, (n + 127) // 128), dims, 'fwd %.1f us  bwd %.1f us' % (a.elapsed_time(b) * 50, b.elapsed_time(c) * 50), flush=True)
