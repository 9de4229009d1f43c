"""Synthetic inputs of the shapes BASELINE.json names: random-init NeRF parameters in the reference's
checkpoint layout, Blender-style rays of one view, and the random tensors render_rays consumes, in the reference's draw
order.  Plain CPU torch; used by bench.py and the tests (the CPU checker re-exports them).  There is no network for datasets
or checkpoints, so every measurement and parity check in this repository runs on these."""
import math
from typing import Dict

import torch

Tensor = torch.Tensor


def init_mlp_params(D=8, W=256, in_xyz=63, in_dir=27, skips=(4,), generator=None) -> Dict[str, Tensor]:
    """Random parameters in the reference checkpoint layout, nn.Linear default init
    (uniform(-1/sqrt(fan_in), 1/sqrt(fan_in)) for weight and bias)."""
    def lin(n_out, n_in):
        bound = 1.0 / math.sqrt(n_in)
        w = (torch.rand(n_out, n_in, generator=generator) * 2 - 1) * bound
        b = (torch.rand(n_out, generator=generator) * 2 - 1) * bound
        return w, b
    p: Dict[str, Tensor] = {}
    for i in range(D):
        n_in = in_xyz if i == 0 else (W + in_xyz if i in skips else W)
        p[f"xyz_encoding_{i + 1}.0.weight"], p[f"xyz_encoding_{i + 1}.0.bias"] = lin(W, n_in)
    p["xyz_encoding_final.weight"], p["xyz_encoding_final.bias"] = lin(W, W)
    p["dir_encoding.0.weight"], p["dir_encoding.0.bias"] = lin(W // 2, W + in_dir)
    p["sigma.weight"], p["sigma.bias"] = lin(1, W)
    p["rgb.0.weight"], p["rgb.0.bias"] = lin(3, W // 2)
    return p


def blender_rays(n: int, seed: int = 1, H: int = 800, W: int = 800, near: float = 2.0, far: float = 6.0,
                 radius: float = 4.0, theta_deg: float = 30.0, phi_deg: float = -30.0) -> Tensor:
    """(n,8) rays of a random pixel subset of one Blender-style view
    (datasets/ray_utils.py:5-24 directions, :27-50 world rays, d normalised :43)."""
    g = torch.Generator().manual_seed(seed)
    focal = 0.5 * W / math.tan(0.5 * 0.6911112070083618)
    pix = torch.randperm(H * W, generator=g)[:n]
    i, j = (pix % W).float(), (pix // W).float()
    dirs = torch.stack([(i - W / 2) / focal, -(j - H / 2) / focal, -torch.ones_like(i)], -1)
    th, ph = math.radians(theta_deg), math.radians(phi_deg)
    t_r = torch.tensor([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, radius], [0, 0, 0, 1.0]])
    r_ph = torch.tensor([[1, 0, 0, 0], [0, math.cos(ph), -math.sin(ph), 0],
                         [0, math.sin(ph), math.cos(ph), 0], [0, 0, 0, 1.0]])
    r_th = torch.tensor([[math.cos(th), 0, -math.sin(th), 0], [0, 1, 0, 0],
                         [math.sin(th), 0, math.cos(th), 0], [0, 0, 0, 1.0]])
    flip = torch.tensor([[-1, 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1.0]])
    c2w = (flip @ r_th @ r_ph @ t_r)[:3, :4]
    rd = dirs @ c2w[:, :3].T
    rd = rd / torch.norm(rd, dim=-1, keepdim=True)
    ro = c2w[:, 3].expand(rd.shape)
    return torch.cat([ro, rd, near * torch.ones_like(ro[:, :1]), far * torch.ones_like(ro[:, :1])], 1).contiguous()


def make_noise(n: int, nc: int, nf: int, seed: int = 2, perturb: bool = True) -> Dict[str, Tensor]:
    """Random tensors in the reference's draw order."""
    g = torch.Generator().manual_seed(seed)
    out: Dict[str, Tensor] = {}
    if perturb:
        out["u_perturb"] = torch.rand(n, nc, generator=g)
    out["noise_coarse"] = torch.randn(n, nc, generator=g)
    if nf > 0:
        out["u_fine"] = torch.rand(n, nf, generator=g)
        out["jitter_fine"] = torch.rand(n, nf, generator=g)
        out["noise_fine"] = torch.randn(n, nc + nf, generator=g)
    return out
