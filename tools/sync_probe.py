"""Host-synchronising calls on the main paths (torch.cuda.set_sync_debug_mode("warn")): a training step with the fused
optimizer, a test-time render, a point query and a depth-only fwd+bwd must queue without making the host wait for the
device.  Finding (round 1): none.      python tools/sync_probe.py"""
import os, sys, warnings
sys.path.insert(0, os.getcwd())
import torch
import __graft_entry__ as ge; ge.build()
import nerf_pl_b200
from nerf_pl_b200.dist import GradBucket
from nerf_pl_b200.losses import loss_dict
from nerf_pl_b200.models import rendering, rendering_shadows as RS
from nerf_pl_b200.models.nerf import Embedding, NeRF
from nerf_pl_b200.optim import FusedAdam
from nerf_pl_b200.datasets.ray_utils import gen_rays
dev = torch.device("cuda:0")
torch.manual_seed(0)
models = [NeRF().to(dev), NeRF().to(dev)]
emb = [Embedding(3, 10), Embedding(3, 4)]
opt = FusedAdam(models, lr=5e-4)
mse = loss_dict["mse"]()
c2w = torch.tensor([[0.6, -0.48, 0.64, 2.56], [0.8, 0.36, -0.48, -1.92], [0.0, 0.8, 0.6, 2.4]])
rays = gen_rays(c2w, 800, 800, 1111.0, 2.0, 6.0, device=dev)[:4096].contiguous()
tgt = torch.rand(4096, 3, device=dev)
def train():
    opt.zero_grad()
    out = rendering.render_rays(models, emb, rays, 64, False, 1.0, 1.0, 64, 1024 * 32, True)
    mse(out, tgt).backward(); opt.step()
def test():
    with torch.no_grad():
        rendering.render_rays(models, emb, rays, 64, False, 0, 0, 64, 1024 * 32, True, True)
def query():
    with torch.no_grad():
        models[1].query(rays[:, :3].contiguous(), rays[:, 3:6].contiguous())
def depth():
    RS.render_rays(models, emb, rays, 64, False, 1.0, 0.0, 64, 1024 * 32, False)["depth_fine"].sum().backward()
for f in (train, test, query, depth):
    for _ in range(2): f()
    torch.cuda.synchronize()
    torch.cuda.set_sync_debug_mode("warn")
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        f()
    torch.cuda.set_sync_debug_mode("default")
    msgs = [str(x.message)[:110] + " @ " + f"{x.filename.split('/')[-1]}:{x.lineno}" for x in w if "synchroniz" in str(x.message).lower()]
    print(f.__name__, "sync warnings:", len(msgs))
    for m in msgs[:6]: print("   ", m)
