"""Golden vectors for the fused Adam step.  Run in the build container:

    python tests/make_golden_adam.py

The optimizer is built the way the reference builds it (utils/__init__.py:17-19: torch.optim.Adam(parameters, lr,
eps=1e-8, weight_decay)); the reference module itself is not imported because utils/__init__.py pulls in its
visualisation dependencies.  Five steps with seeded gradients, two cases (weight_decay 0 and 1e-4), fp32 on the CPU."""
import os

import numpy as np
import torch

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "adam.npz")


def run(weight_decay, seed, n=4099, steps=5, lr=5e-4):
    g = torch.Generator().manual_seed(seed)
    p = torch.nn.Parameter(torch.randn(n, generator=g) * 0.1)
    p0 = p.detach().clone()
    opt = torch.optim.Adam([p], lr=lr, eps=1e-8, weight_decay=weight_decay)
    grads, params = [], []
    for _ in range(steps):
        gr = torch.randn(n, generator=g) * torch.logspace(-6, 0, n)      # gradients across six decades
        p.grad = gr.clone()
        opt.step()
        grads.append(gr.numpy().copy())
        params.append(p.detach().numpy().copy())
    st = opt.state[p]
    return dict(p0=p0.numpy(), grads=np.stack(grads), params=np.stack(params), exp_avg=st["exp_avg"].numpy(),
                exp_avg_sq=st["exp_avg_sq"].numpy(), lr=np.float64(lr), weight_decay=np.float64(weight_decay))


if __name__ == "__main__":
    out = {}
    for name, wd, seed in (("wd0", 0.0, 11), ("wd1e-4", 1e-4, 12)):
        for k, v in run(wd, seed).items():
            out[f"{name}.{k}"] = v
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")
