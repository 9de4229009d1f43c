#!/usr/bin/env python
"""nrf_peer_allreduce_mean (csrc/peer_allreduce.cu) against NCCL on N GPUs of one node:
   torchrun --nproc-per-node N --master-addr 127.0.0.1 tools/peer_allreduce_check.py
values against NCCL's sum / N, bit-identity across the ranks, 100 back-to-back rounds, a CUDA-graph replay, timing."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from nerf_pl_b200.dist import PeerExchange

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
n = 2 * 595844
px = PeerExchange(n, dev)
if not px.ok:
    if rank == 0:
        print(json.dumps({"peer_allreduce": "unavailable", "error": getattr(px, "error", "")}))
    dist.destroy_process_group()
    sys.exit(0)
g = torch.Generator(device=dev).manual_seed(100 + rank)
worst, ident = 0.0, True
for it in range(100):
    x = torch.randn(n, device=dev, generator=g) * (1 + it % 7)
    ref = x.clone()
    dist.all_reduce(ref)
    ref /= world
    px.flat.copy_(x)
    px.allreduce_mean()
    worst = max(worst, float((px.flat - ref).abs().max() / ref.abs().max()))
    lo, hi = px.flat.clone(), px.flat.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    ident = ident and bool(torch.equal(lo, hi))
# CUDA graph
px.flat.copy_(torch.randn(n, device=dev, generator=g))
torch.cuda.synchronize()
dist.barrier()
graph = torch.cuda.CUDAGraph()
with torch.cuda.graph(graph):
    px.allreduce_mean()
graph_ok = True
for it in range(20):
    x = torch.randn(n, device=dev, generator=g)
    ref = x.clone()
    dist.all_reduce(ref)
    ref /= world
    px.flat.copy_(x)
    graph.replay()
    graph_ok = graph_ok and float((px.flat - ref).abs().max() / ref.abs().max()) < 1e-5


def timed(fn, reps=200):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps * 1e3], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


y = torch.randn(n, device=dev)


def nccl():
    y.div_(world)
    dist.all_reduce(y)


t_peer, t_nccl = timed(px.allreduce_mean), timed(nccl)
if rank == 0:
    print(json.dumps({"world": world, "n_floats": n, "max_rel_diff_vs_nccl": worst, "bit_identical_across_ranks": ident,
                      "graph_replay_ok": graph_ok, "us_per_call_peer_kernel": t_peer, "us_per_call_div_plus_nccl": t_nccl}))
del graph
px.close()
dist.barrier()
dist.destroy_process_group()
