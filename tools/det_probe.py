"""Does a test-time render depend on how the rays are batched?  Renders 65 536 rays whole, twice, shifted by 1000 rays,
as the second half only and in chunks of 8192, and compares bit for bit.  Finding: opacity_coarse is bit-identical in every
case; the fine keys differ between ANY two calls because the fork's sample_pdf draws fresh uniforms on every call, test
time included (rendering.py:36-41) - inject u_fine / jitter_fine (tools/frame_bench.py, tests/test_gpu_edges.py) and the
fine keys are bit-identical too.      python tools/det_probe.py"""
import math, os, sys
sys.path.insert(0, os.getcwd())
import torch
import __graft_entry__ as ge; ge.build()
import nerf_pl_b200
from nerf_pl_b200.datasets.ray_utils import gen_rays
from nerf_pl_b200.models import rendering
from nerf_pl_b200.models.nerf import Embedding, NeRF
dev=torch.device("cuda:0")
torch.manual_seed(0)
models=[NeRF().to(dev), NeRF().to(dev)]
emb=[Embedding(3,10), Embedding(3,4)]
g=torch.Generator().manual_seed(1)
o=torch.randn(65536,3,generator=g); o=4*o/o.norm(dim=1,keepdim=True)
d=torch.nn.functional.normalize(-o+0.3*torch.randn(65536,3,generator=g),dim=1)
rays=torch.cat([o,d,torch.full((65536,1),2.0),torch.full((65536,1),6.0)],1).to(dev)
def R(r, chunk=32768):
    with torch.no_grad():
        outs=[rendering.render_rays(models,emb,r[i:i+chunk],64,False,0,0,64,1024*32,True,True) for i in range(0,r.shape[0],chunk)]
        return {k: torch.cat([x[k] for x in outs],0) for k in outs[0]}
a=R(rays); b=R(rays)
for k in a: print("repeat", k, torch.equal(a[k],b[k]), torch.isnan(a[k]).sum().item(), (a[k]-b[k]).abs().max().item())
c=R(rays[1000:])   # shifted chunk boundaries
for k in a: print("shift", k, torch.equal(a[k][1000:],c[k]), (a[k][1000:]-c[k]).abs().max().item(), (a[k][1000:]!=c[k]).sum().item())
c=R(rays[32768:])
for k in a: print("half", k, torch.equal(a[k][32768:],c[k]), (a[k][32768:]!=c[k]).sum().item())
c=R(rays, chunk=8192)
for k in a: print("chunk8192", k, torch.equal(a[k],c[k]), (a[k]!=c[k]).sum().item())
