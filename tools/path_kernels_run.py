"""One launch of each HBM-bound path kernel at a 32768-ray test-time chunk (the sizes of bench.py's
side_rows.path_kernels_32768_rays), for an ncu capture:
    ncu --set full --clock-control none -k regex:"composite|sample_|merge_sort" --launch-skip 10 -c 5 python tools/path_kernels_run.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import __graft_entry__ as ge
ge.build()
from nerf_pl_b200 import ops
from nerf_pl_b200.datasets.ray_utils import gen_rays

dev = torch.device("cuda:0")
NC = NF = 64
Nr = 32768
c2w = torch.tensor([[0.6, -0.48, 0.64, 2.56], [0.8, 0.36, -0.48, -1.92], [0.0, 0.8, 0.6, 2.4]])
rays = gen_rays(c2w, 800, 800, 1111.0, 2.0, 6.0, device=dev)[:Nr].contiguous()
u, j = torch.rand(Nr, NC, device=dev), torch.rand(Nr, NF, device=dev)
w_c = torch.rand(Nr, NC, device=dev)
S = NC + NF
sig, col, nz = torch.randn(Nr, S, device=dev), torch.rand(Nr, S, 3, device=dev), torch.randn(Nr, S, device=dev)
g3, g1 = torch.rand(Nr, 3, device=dev), torch.rand(Nr, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for rep in range(3):       # 2 warm-up rounds (10 launches to skip), then the captured one
    z_c = ops.sample_coarse(rays, NC, False, 1.0, u)
    z_f = ops.sample_pdf(rays, w_c, NF, j, j)
    z_a = ops.merge_sort(z_c, z_f)
    flush.zero_()
    ops.composite(sig, col, z_a, rays, nz, 1.0, True)
    flush.zero_()
    ops.composite_bwd(sig, col, z_a, rays, nz, 1.0, True, g3, g1, g1, None)
torch.cuda.synchronize()
print("done")
