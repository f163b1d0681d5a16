"""Targeted launch of the fused-MLP kernels at the NGP step's size, for ncu captures."""
import sys
import torch
sys.path.insert(0, '.')
from arcnerf_b200 import functional as F

n = int(sys.argv[1]) if len(sys.argv) > 1 else 147624
dev = torch.device('cuda:0')
for dims, acts in (((32, 64, 1, 16), ('relu', 'none')), ((32, 64, 2, 3), ('relu', 'sigmoid'))):
    spec = F.MlpSpec(*dims, *acts)
    params = (torch.rand(spec.n_params, device=dev) - 0.5) * 0.5
    blob = F.mlp_pack(params, spec)
    x = (torch.rand((n, 32), device=dev) - 0.5).half()
    dy = torch.randn((n, dims[3]), device=dev) * 1e-2
    out = torch.empty((n, dims[3]), dtype=torch.float32, device=dev)
    dW = torch.zeros(spec.n_params, device=dev)
    dx = torch.empty((n, 32), dtype=torch.float32, device=dev)
    for _ in range(3):
        F.mlp_forward(x, blob, spec, out=out)
        F.mlp_backward(x, blob, spec, dy, dW=dW, dx=dx)
    torch.cuda.synchronize()
    a, b, c = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    a.record()
    for _ in range(10):
        F.mlp_forward(x, blob, spec, out=out)
    b.record()
    for _ in range(10):
        F.mlp_backward(x, blob, spec, dy, dW=dW, dx=dx)
    c.record()
    torch.cuda.synchronize()
    print(dims, 'fwd %.1f us  bwd %.1f us' % (a.elapsed_time(b) * 100, b.elapsed_time(c) * 100))
