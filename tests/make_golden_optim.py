"""Golden vectors for the fused sgd / radam / ranger steps (utils/__init__.py:14-27).  Run in the
build container (needs /root/reference):

    python tests/make_golden_optim.py

RAdam and Ranger are the reference's own classes (utils/optimizers.py, imported by path: utils/__init__.py would pull in
the visualisation dependencies), built the way get_optimizer builds them (lr, eps=1e-8, weight_decay); SGD is
torch.optim.SGD(lr, momentum, weight_decay).  Eight steps with seeded gradients - past the N_sma >= 5 switch to the
rectified step (step 6 for beta2 = 0.999) and Ranger's first lookahead (k = 6) -, fp32 on the CPU."""
import importlib.util
import os
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "golden", "optim.npz")
REF = os.environ.get("NERF_REFERENCE", "/root/reference")


def ref_optimizers():
    spec = importlib.util.spec_from_file_location("ref_optimizers", os.path.join(REF, "utils", "optimizers.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def run(make, seed, n=1031, steps=8):
    g = torch.Generator().manual_seed(seed)
    p = torch.nn.Parameter(torch.randn(n, generator=g) * 0.1)
    p0 = p.detach().clone()
    opt = make([p])
    grads, params = [], []
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for _ in range(steps):
            gr = torch.randn(n, generator=g) * torch.logspace(-6, 0, n)      # gradients across six decades
            p.grad = gr.clone()
            opt.step()
            grads.append(gr.numpy().copy())
            params.append(p.detach().numpy().copy())
    out = dict(p0=p0.numpy(), grads=np.stack(grads), params=np.stack(params))
    for k, v in opt.state[p].items():
        if torch.is_tensor(v):
            out[k] = v.numpy()
    return out


if __name__ == "__main__":
    R = ref_optimizers()
    lr = 5e-4
    cases = {
        "sgd_plain": lambda ps: torch.optim.SGD(ps, lr=lr, momentum=0.0, weight_decay=0.0),
        "sgd_mom_wd": lambda ps: torch.optim.SGD(ps, lr=lr, momentum=0.9, weight_decay=1e-4),
        "radam_wd0": lambda ps: R.RAdam(ps, lr=lr, eps=1e-8, weight_decay=0),
        "radam_wd": lambda ps: R.RAdam(ps, lr=lr, eps=1e-8, weight_decay=1e-4),
        "radam_nosgd": lambda ps: R.RAdam(ps, lr=lr, eps=1e-8, weight_decay=1e-4, degenerated_to_sgd=False),
        "ranger_wd0": lambda ps: R.Ranger(ps, lr=lr, eps=1e-8, weight_decay=0),
        "ranger_wd": lambda ps: R.Ranger(ps, lr=lr, eps=1e-8, weight_decay=1e-4),
        "ranger_k2": lambda ps: R.Ranger(ps, lr=lr, alpha=0.3, k=2, eps=1e-8, weight_decay=0),
    }
    out = {"lr": np.float64(lr)}
    for i, (name, make) in enumerate(cases.items()):
        for k, v in run(make, 20 + i).items():
            out[f"{name}.{k}"] = v
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")
