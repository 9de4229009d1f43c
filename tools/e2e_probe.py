"""Where does the end-to-end step time go?  Times the graphed training step of bench.py (config 2) in five ways over K
steps each: event-timed replays with an L2 flush and idle gaps (bench `value`), back-to-back replays (one event pair
around all K), back-to-back with the H2D copies, the pipelined submit() loop (wall clock) and the blocking loop.
Samples SM clock and power while each loop runs.      python tools/e2e_probe.py [K]"""
import os
import subprocess
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


class Smi:
    def __init__(self):
        self.rows, self.stop_flag = [], False

    def start(self):
        self.rows, self.stop_flag = [], False
        self.t = threading.Thread(target=self.run, daemon=True)
        self.t.start()

    def run(self):
        while not self.stop_flag:
            try:
                o = subprocess.run(["nvidia-smi", "-i", "0", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                self.rows.append((float(o[0]), float(o[1])))
            except Exception:
                pass

    def stop(self):
        self.stop_flag = True
        self.t.join()
        if not self.rows:
            return None
        c = sorted(r[0] for r in self.rows)
        p = sorted(r[1] for r in self.rows)
        return {"sm_mhz_med": c[len(c) // 2], "sm_mhz_min": c[0], "power_w_med": p[len(p) // 2], "n": len(c)}


def main():
    K = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    import __graft_entry__ as ge
    ge.build()
    import nerf_pl_b200
    from nerf_pl_b200.dist import GradBucket
    from nerf_pl_b200.graph import GraphedStep
    from nerf_pl_b200.losses import loss_dict
    from nerf_pl_b200.models.nerf import Embedding, NeRF
    from nerf_pl_b200.models.rendering import render_rays
    from nerf_pl_b200.optim import FusedAdam
    dev = torch.device("cuda:0")
    nerf_pl_b200.set_mlp_precision("bf16")
    torch.manual_seed(0)
    models = [NeRF().to(dev), NeRF().to(dev)]
    emb = [Embedding(3, 10), Embedding(3, 4)]
    bucket = GradBucket(models)
    opt = FusedAdam(models, lr=5e-4, bucket=bucket, capturable=True)
    mse = loss_dict["mse"]()
    B = 4096
    g = torch.Generator().manual_seed(1)
    o = torch.randn(B, 3, generator=g)
    o = 4.0 * o / o.norm(dim=1, keepdim=True)
    d = torch.nn.functional.normalize(-o + 0.3 * torch.randn(B, 3, generator=g), dim=1)
    rays_host = torch.cat([o, d, torch.full((B, 1), 2.0), torch.full((B, 1), 6.0)], 1).pin_memory()
    rgbs_host = torch.rand(B, 3, generator=g).pin_memory()
    rays_dev, rgbs_dev = rays_host.to(dev), rgbs_host.to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def step(r, t):
        opt.zero_grad()
        out = render_rays(models, emb, r, 64, False, 1.0, 1.0, 64, 1024 * 32, True)
        loss = mse(out, t)
        loss.backward()
        bucket.allreduce_mean()
        opt.step()
        return loss

    gs = GraphedStep(step, (rays_dev, rgbs_dev), warmup=3, optimizers=[opt])
    for _ in range(5):
        gs(rays_host, rgbs_host)
    torch.cuda.synchronize()
    smi = Smi()
    E = lambda: torch.cuda.Event(enable_timing=True)

    def report(name, ms, extra=None):
        print(f"{name:46s} {ms:7.4f} ms/step  {extra or ''}", flush=True)

    # 1 bench `value`: flush + per-step event pair
    ev = [(E(), E()) for _ in range(K)]
    smi.start()
    for i in range(K):
        flush.zero_()
        ev[i][0].record()
        gs(*gs.static_in)
        ev[i][1].record()
    torch.cuda.synchronize()
    report("flush + per-step events (bench value)", sum(a.elapsed_time(b) for a, b in ev) / K, smi.stop())
    # 2 back to back, no flush
    a, b = E(), E()
    smi.start()
    a.record()
    for i in range(K):
        gs(*gs.static_in)
    b.record()
    torch.cuda.synchronize()
    report("back-to-back replays, one event pair", a.elapsed_time(b) / K, smi.stop())
    # 3 back to back with H2D
    smi.start()
    a.record()
    for i in range(K):
        gs(rays_host, rgbs_host)
    b.record()
    torch.cuda.synchronize()
    report("back-to-back replays + H2D copies", a.elapsed_time(b) / K, smi.stop())
    # 4 pipelined submit
    smi.start()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    pend = None
    for i in range(K):
        nxt = gs.submit(rays_host, rgbs_host)
        if pend is not None:
            pend.result()
        pend = nxt
    pend.result()
    report("pipelined submit() loop, wall clock", (time.perf_counter() - t0) / K * 1e3, smi.stop())
    # 5 blocking
    smi.start()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(K):
        gs(rays_host, rgbs_host).item()
    report("blocking loop, wall clock", (time.perf_counter() - t0) / K * 1e3, smi.stop())
    # 6 host cost of one submit (no GPU wait): queue K submits without collecting
    t0 = time.perf_counter()
    for i in range(8):
        gs(rays_host, rgbs_host)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    report("host time to queue one step", (t1 - t0) / 8 * 1e3)
    gs.close()


if __name__ == "__main__":
    main()
