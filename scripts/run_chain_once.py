import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from arcnerf_b200.plugin.nerf_mlp import FusedNeRFNets
nets = FusedNeRFNets()
n = 196608
x = torch.rand((n, 3), device='cuda') * 2 - 1; d = torch.randn((n, 3), device='cuda')
with torch.no_grad():
    for _ in range(3):
        nets(x, d)
torch.cuda.synchronize()
