"""Golden vectors for the rows either side of the path from the REFERENCE itself (build container
only): datasets/ray_utils.py (get_ray_directions / get_rays / get_ndc_rays), losses.py (MSELoss, SMMSELoss) and
metrics.py (mse / psnr), imported unmodified from /root/reference.  kornia is absent: `create_meshgrid` (the only kornia
call on this path, ray_utils.py:17) is shimmed with its documented behaviour (x = 0..W-1 over columns, y = 0..H-1 over
rows, shape (1,H,W,2)); kornia.losses.ssim is shimmed with a stub (ssim is not used).

Run:  python tests/make_golden_rays.py      -> tests/golden/rays_loss.npz"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("NERF_REFERENCE", "/root/reference")


def import_reference():
    kornia = types.ModuleType("kornia")

    def create_meshgrid(height, width, normalized_coordinates=True):
        assert not normalized_coordinates
        xs = torch.linspace(0, width - 1, width)
        ys = torch.linspace(0, height - 1, height)
        return torch.stack(torch.meshgrid([xs, ys], indexing="ij"), -1).permute(1, 0, 2).unsqueeze(0)   # kornia 0.2.0 grid.py

    kornia.create_meshgrid = create_meshgrid
    kl = types.ModuleType("kornia.losses")
    kl.ssim = lambda *a, **k: (_ for _ in ()).throw(NotImplementedError)
    kornia.losses = kl
    sys.modules["kornia"], sys.modules["kornia.losses"] = kornia, kl
    sys.path.insert(0, REF)
    import importlib.util

    def load(name, rel):
        spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod
    return load("ref_ray_utils", "datasets/ray_utils.py"), load("ref_losses", "losses.py"), load("ref_metrics", "metrics.py")


def pose(seed):
    g = torch.Generator().manual_seed(seed)
    a = torch.randn(3, 3, generator=g)
    q, _ = torch.linalg.qr(a)
    return torch.cat([q, torch.randn(3, 1, generator=g) * 2 + torch.tensor([[0.5], [-0.3], [4.0]])], 1).contiguous()


def forward_facing_pose(seed):
    """LLFF-like: the camera looks down -z with a small rotation and offset, so NDC is well conditioned."""
    g = torch.Generator().manual_seed(seed)
    ang = (torch.rand(3, generator=g) - 0.5) * 0.4
    cx, cy, cz = torch.cos(ang)
    sx, sy, sz = torch.sin(ang)
    rx = torch.tensor([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    ry = torch.tensor([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    rz = torch.tensor([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return torch.cat([rz @ ry @ rx, (torch.rand(3, 1, generator=g) - 0.5) * 0.6], 1).float().contiguous()


def main():
    ru, losses, metrics = import_reference()
    rec = {}
    cases = [("blender_small", 20, 24, 30.5, 2.0, 6.0, False, 1), ("fork_bounds", 16, 16, 22.0, 1.0, 200.0, False, 2),
             ("llff_ndc", 18, 24, 19.4, 1.0, 1.0, True, 3)]
    for name, H, W, focal, near, far, ndc, seed in cases:
        c2w = forward_facing_pose(seed) if ndc else pose(seed)
        d = ru.get_ray_directions(H, W, focal)
        o, dd = ru.get_rays(d, c2w)
        rec[f"{name}/hwf"] = np.array([H, W, focal, near, far, float(ndc)], dtype=np.float64)
        rec[f"{name}/c2w"] = c2w.numpy()
        rec[f"{name}/directions"] = d.numpy()
        rec[f"{name}/rays_o"], rec[f"{name}/rays_d"] = o.numpy(), dd.numpy()
        if ndc:
            o, dd = ru.get_ndc_rays(H, W, focal, near, o, dd)
            rec[f"{name}/ndc_o"], rec[f"{name}/ndc_d"] = o.numpy(), dd.numpy()
            nr, fr = 0.0, 1.0
        else:
            nr, fr = near, far
        rec[f"{name}/rays"] = torch.cat([o, dd, nr * torch.ones_like(o[:, :1]), fr * torch.ones_like(o[:, :1])], 1).numpy()
    g = torch.Generator().manual_seed(7)
    for name, n in (("loss_small", 37), ("loss_batch", 4096)):
        a = torch.rand(n, 3, generator=g, requires_grad=True)
        b = torch.rand(n, 3, generator=g, requires_grad=True)
        t = torch.rand(n, 3, generator=g)
        l2 = losses.MSELoss()({"rgb_coarse": a, "rgb_fine": b}, t)
        l2.backward()
        rec[f"{name}/a"], rec[f"{name}/b"], rec[f"{name}/t"] = a.detach().numpy(), b.detach().numpy(), t.numpy()
        rec[f"{name}/mse2"], rec[f"{name}/ga"], rec[f"{name}/gb"] = l2.detach().numpy(), a.grad.numpy(), b.grad.numpy()
        rec[f"{name}/mse1"] = losses.MSELoss()({"rgb_coarse": a.detach()}, t).numpy()
        rec[f"{name}/sm2"] = losses.SMMSELoss()({"sm_coarse": a.detach(), "sm_fine": b.detach()}, t).numpy()
        mask = torch.rand(n, 3, generator=g) > 0.4
        rec[f"{name}/mask"] = mask.numpy()
        rec[f"{name}/psnr"] = metrics.psnr(a.detach(), t).numpy()
        rec[f"{name}/psnr_masked"] = metrics.psnr(a.detach(), t, mask).numpy()
        rec[f"{name}/mse_masked"] = metrics.mse(a.detach(), t, mask).numpy()
    out = os.path.join(HERE, "golden", "rays_loss.npz")
    np.savez_compressed(out, **rec)
    print(out, os.path.getsize(out))


if __name__ == "__main__":
    main()
