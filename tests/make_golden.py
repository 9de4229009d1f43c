"""Generate tests/golden/*.npz by running the REFERENCE itself (build container only).

Imports /root/reference/models/*.py unmodified (a `torchsearchsorted` shim forwards to
torch.searchsorted, which SURVEY §8c verified identical to the reference extension; the
extension itself is additionally compiled into oracle/_ref by oracle/build_ref.py and
compared in tests/test_oracle_golden.py), injects the random tensors by patching
torch.rand / torch.rand_like / torch.randn, and stores inputs + outputs + gradients.

Run:  python tests/make_golden.py            (needs /root/reference; not needed on the GPU box)
"""
import contextlib
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("NERF_REFERENCE", "/root/reference")
sys.path.insert(0, ROOT)

from oracle import nerf_oracle as O  # noqa: E402  (used only for input generators / param init)


def import_reference():
    shim = types.ModuleType("torchsearchsorted")
    shim.searchsorted = lambda a, v, out=None, side="left": torch.searchsorted(
        a.contiguous(), v.contiguous(), right=(side == "right"))
    sys.modules["torchsearchsorted"] = shim
    sys.path.insert(0, REF)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        import models.nerf as nerf
        import models.rendering as rendering
        import models.rendering_shadows as rendering_shadows
        import models.rendering_rgb_sm as rendering_rgb_sm
        import models.efficient_shadow_mapping as esm
        import models.camera as camera
    return nerf, rendering, rendering_shadows, rendering_rgb_sm, esm, camera


@contextlib.contextmanager
def injected_rng(tensors):
    """Make torch.rand/rand_like/randn return the given tensors in order."""
    queue = list(tensors)
    orig = (torch.rand, torch.rand_like, torch.randn)

    def pop(shape):
        t = queue.pop(0)
        assert tuple(t.shape) == tuple(shape), (t.shape, shape)
        return t.clone()

    def _shape(a):
        return tuple(a[0]) if len(a) == 1 and isinstance(a[0], (tuple, list, torch.Size)) else tuple(a)

    torch.rand = lambda *a, **k: pop(_shape(a))
    torch.rand_like = lambda x, **k: pop(x.shape)
    torch.randn = lambda *a, **k: pop(_shape(a))
    try:
        yield
        assert not queue, f"{len(queue)} injected tensors unused"
    finally:
        torch.rand, torch.rand_like, torch.randn = orig


def ref_models(nerf, W, seeds=(10, 11), D=8, skips=(4,)):
    params, models = [], []
    for s in seeds:
        p = O.init_mlp_params(D=D, W=W, skips=skips, generator=torch.Generator().manual_seed(s))
        m = nerf.NeRF(D=D, W=W, skips=list(skips))
        m.load_state_dict(p, strict=True)      # also pins the checkpoint key/shape layout
        params.append(p)
        models.append(m)
    return params, models


def noise_list(noise, perturb, nf):
    seq = []
    if perturb > 0:
        seq.append(noise["u_perturb"])
    seq.append(noise["noise_coarse"])
    if nf > 0:
        seq += [noise["u_fine"], noise["jitter_fine"], noise["noise_fine"]]
    return seq


def run_render_case(mod, nerf, name, *, variant, N, Nc, Nf, W, use_disp, perturb, noise_std,
                    white_back, test_time, near, far, ray_seed, out):
    params, models = ref_models(nerf, W)
    emb = [nerf.Embedding(3, 10), nerf.Embedding(3, 4)]
    rays = O.blender_rays(N, seed=ray_seed, near=near, far=far)
    noise = O.make_noise(N, Nc, Nf, seed=ray_seed + 100, perturb=perturb > 0)
    with injected_rng(noise_list(noise, perturb, Nf)):
        res = mod.render_rays(models, emb, rays, Nc, use_disp, perturb, noise_std, Nf, 1024 * 32,
                              white_back, test_time)
    keys = sorted(res.keys())
    g = torch.Generator().manual_seed(ray_seed + 200)
    probe = {k: torch.rand(res[k].shape, generator=g) for k in keys}
    loss = sum((res[k] * probe[k]).sum() for k in keys)
    loss.backward()
    rec = {"rays": rays.numpy(), "cfg": np.array([N, Nc, Nf, W, int(use_disp), int(white_back), int(test_time)]),
           "cfgf": np.array([perturb, noise_std, near, far], dtype=np.float64), "keys": np.array(keys)}
    for k, v in noise.items():
        rec["noise_" + k] = v.numpy()
    for k in keys:
        rec["out_" + k] = res[k].detach().numpy()
        rec["probe_" + k] = probe[k].numpy()
    for mi, m in enumerate(models):
        for pn, pv in m.named_parameters():
            rec[f"grad{mi}_{pn}"] = (pv.grad if pv.grad is not None else torch.zeros_like(pv)).numpy()
    out[name] = rec
    print(name, {k: tuple(res[k].shape) for k in keys})


def main():
    torch.set_num_threads(4)
    nerf, rendering, rendering_shadows, rendering_rgb_sm, esm, camera = import_reference()
    gold = {}

    # --- render_rays, the three variants -------------------------------------------------
    run_render_case(rendering, nerf, "rgb_train", variant="rgb", N=24, Nc=16, Nf=16, W=64, use_disp=False,
                    perturb=1.0, noise_std=1.0, white_back=True, test_time=False, near=2.0, far=6.0,
                    ray_seed=1, out=gold)
    run_render_case(rendering, nerf, "rgb_test", variant="rgb", N=24, Nc=16, Nf=24, W=64, use_disp=False,
                    perturb=0.0, noise_std=0.0, white_back=True, test_time=True, near=2.0, far=6.0,
                    ray_seed=2, out=gold)
    run_render_case(rendering, nerf, "rgb_disp_fork_bounds", variant="rgb", N=16, Nc=16, Nf=0, W=64,
                    use_disp=True, perturb=0.5, noise_std=0.0, white_back=False, test_time=False,
                    near=1.0, far=200.0, ray_seed=3, out=gold)
    run_render_case(rendering_shadows, nerf, "depth_train", variant="depth", N=24, Nc=16, Nf=16, W=64,
                    use_disp=False, perturb=1.0, noise_std=1.0, white_back=True, test_time=False,
                    near=1.0, far=200.0, ray_seed=4, out=gold)
    run_render_case(rendering_rgb_sm, nerf, "rgb_sm_train", variant="rgb_sm", N=24, Nc=16, Nf=16, W=64,
                    use_disp=False, perturb=1.0, noise_std=0.0, white_back=False, test_time=False,
                    near=2.0, far=6.0, ray_seed=5, out=gold)
    np.savez_compressed(os.path.join(HERE, "golden", "render_rays.npz"),
                        **{f"{c}/{k}": v for c, r in gold.items() for k, v in r.items()})

    # --- Embedding / NeRF (full width, few rows) ------------------------------------------
    g = torch.Generator().manual_seed(7)
    x3 = (torch.rand(16, 3, generator=g) * 2 - 1) * 6.0
    x3[0] = torch.tensor([200.0, -150.0, 99.5])           # fork bounds: |512 x| ~ 1e5 rad
    rec = {"x3": x3.numpy(), "emb10": nerf.Embedding(3, 10)(x3).numpy(), "emb4": nerf.Embedding(3, 4)(x3).numpy(),
           "emb4_lin": nerf.Embedding(3, 4, logscale=False)(x3).numpy()}
    p256 = O.init_mlp_params(generator=torch.Generator().manual_seed(20))
    m256 = nerf.NeRF()
    m256.load_state_dict(p256, strict=True)
    x90 = torch.cat([nerf.Embedding(3, 10)(x3), nerf.Embedding(3, 4)(x3 / x3.norm(dim=-1, keepdim=True))], -1)
    rec["x90"] = x90.numpy()
    rec["nerf256_full"] = m256(x90).detach().numpy()
    rec["nerf256_sigma"] = m256(x90[:, :63], sigma_only=True).detach().numpy()
    rec["nerf256_wsum"] = np.array([float(sum(v.double().sum() for v in p256.values()))])
    rec["nerf_keys"] = np.array(list(m256.state_dict().keys()))
    rec["nerf_shapes"] = np.array([str(tuple(v.shape)) for v in m256.state_dict().values()])
    np.savez_compressed(os.path.join(HERE, "golden", "nerf_modules.npz"), **rec)

    # --- sample_pdf alone (cdf, indices) ---------------------------------------------------
    g = torch.Generator().manual_seed(8)
    N, nb, nf = 64, 62, 64
    w = torch.rand(N, nb, generator=g) ** 4
    w[:4] = 0.0                                            # all-eps rows
    w[4, 10] = 50.0                                        # one dominant bin
    rays = O.blender_rays(N, seed=8, near=2.0, far=6.0)
    u = torch.rand(N, nf, generator=g)
    u[5, :4] = torch.tensor([0.0, 1.0 - 2 ** -24, 0.5, 0.25])
    jit = torch.rand(N, nf, generator=g)
    with injected_rng([u, jit]):
        z = rendering.sample_pdf(rays, w, nf)
    with injected_rng([u, jit]):
        z_sh = rendering_shadows.sample_pdf(rays, w, nf)
    assert torch.equal(z, z_sh)
    np.savez_compressed(os.path.join(HERE, "golden", "sample_pdf.npz"), rays=rays.numpy(), w=w.numpy(),
                        u=u.numpy(), jitter=jit.numpy(), z=z.numpy())

    # --- searchsorted known answers (torchsearchsorted/test/test_searchsorted.py:9-44) -----
    sys.path.insert(0, os.path.join(REF, "torchsearchsorted", "src", "torchsearchsorted"))
    import utils as tss_utils                               # numpy_searchsorted, the test's own oracle
    rng = np.random.RandomState(9)
    ss = {}
    ci = 0
    for (Ba, Bv, A, V) in [(1, 1, 1, 1), (100, 100, 50, 12), (1, 100, 50, 12), (100, 1, 50, 120),
                           (200, 200, 500, 120), (100, 100, 1, 12), (64, 64, 63, 64)]:
        a = np.sort(rng.rand(Ba, A).astype(np.float32), axis=1)
        v = rng.rand(Bv, V).astype(np.float32)
        if A > 3:
            v[0, : min(V, 3)] = a[0, : min(V, 3)]           # exact ties
        ss[f"c{ci}_a"], ss[f"c{ci}_v"] = a, v
        for side in ("left", "right"):
            ss[f"c{ci}_{side}"] = tss_utils.numpy_searchsorted(a, v, side=side).astype(np.int64)
        ci += 1
    ss["tie_a"] = np.array([[0, .25, .25, .5, 1]], dtype=np.float32)
    ss["tie_v"] = np.array([[0, .25, .5, 1, 1.5, -1]], dtype=np.float32)
    ss["tie_right"] = np.array([[1, 3, 4, 5, 5, 0]], dtype=np.int64)      # SURVEY §8c, measured on the built extension
    ss["tie_left"] = np.array([[0, 1, 3, 4, 5, 0]], dtype=np.int64)
    ss["n_cases"] = np.array(ci)
    np.savez_compressed(os.path.join(HERE, "golden", "searchsorted.npz"), **ss)

    # --- shadow mapping chain (efficient_sm) -----------------------------------------------
    H = Wd = 16
    res = (Wd, H)
    g = torch.Generator().manual_seed(12)

    def lookat_c2w(eye):
        return torch.from_numpy(camera.Camera.c2w_from_lookat(np.array(eye, dtype=np.float32),
                                                              np.zeros(3, dtype=np.float32)))[:3, :4]

    cams = []
    for eye in ([0.0, 1.0, 4.0], [2.0, 1.5, 3.0]):
        c = camera.Camera(30, res)
        c.set_pose_using_blender_matrix(lookat_c2w(eye))
        cams.append(c)
    light = camera.Camera(30, res)
    light.set_pose_using_blender_matrix(lookat_c2w([1.0, 4.0, 2.0]))
    n = 40
    seg = torch.tensor([0] * 25 + [1] * 15)
    ppc = {"eye_pos": torch.stack([cams[s].eye_pos for s in seg]),
           "camera": torch.stack([cams[s].camera for s in seg])}
    ii, jj = torch.meshgrid(torch.arange(Wd).float(), torch.arange(H).float(), indexing="xy")
    light_pixels = torch.stack([ii.reshape(-1) + .5, jj.reshape(-1) + .5, torch.ones(H * Wd)], 1)
    sel = torch.randperm(H * Wd, generator=g)[:n]
    cam_pixels = light_pixels[sel].clone()
    rec = {"res": np.array(res), "cam_pixels": cam_pixels.numpy(), "light_pixels": light_pixels.numpy(),
           "eye_pos": ppc["eye_pos"].numpy(), "camera": ppc["camera"].numpy(),
           "light_eye": light.eye_pos.numpy(), "light_camera": light.camera.numpy()}
    for method in ("shadow_method_1", "shadow_method_2"):
        for lni in (0, 8):
            cd = {k: (3.0 + 2.0 * torch.rand(n, generator=g)).requires_grad_() for k in ("depth_coarse", "depth_fine")}
            ld = {k: (3.0 + 2.0 * torch.rand(H * Wd, generator=g)).requires_grad_() for k in ("depth_coarse", "depth_fine")}
            out = rendering_shadows.efficient_sm(cam_pixels, light_pixels, dict(cd), dict(ld), ppc, light, res,
                                                 True, lni, method)
            out2 = rendering_rgb_sm.efficient_sm(cam_pixels, light_pixels, dict(cd), dict(ld), ppc, light, res,
                                                 True, lni, method)
            assert torch.equal(out["rgb_coarse"], out2["sm_coarse"]) and torch.equal(out["rgb_fine"], out2["sm_fine"])
            probe = torch.rand(n, 3, generator=g)
            ((out["rgb_coarse"] * probe).sum() + 2 * (out["rgb_fine"] * probe).sum()).backward()
            tag = f"{method}_lni{lni}"
            for k in cd:
                rec[f"{tag}/cam_{k}"] = cd[k].detach().numpy()
                rec[f"{tag}/light_{k}"] = ld[k].detach().numpy()
                rec[f"{tag}/gcam_{k}"] = (cd[k].grad if cd[k].grad is not None else torch.zeros_like(cd[k])).numpy()
                rec[f"{tag}/glight_{k}"] = (ld[k].grad if ld[k].grad is not None else torch.zeros_like(ld[k])).numpy()
            rec[f"{tag}/rgb_coarse"] = out["rgb_coarse"].detach().numpy()
            rec[f"{tag}/rgb_fine"] = out["rgb_fine"].detach().numpy()
            rec[f"{tag}/probe"] = probe.numpy()
    # Camera algebra (models/camera.py:20-31, 69-89, 121-132)
    R, Q = cams[0].get_transformation_to(light)
    rec["cam0_M"], rec["cam0_eye"] = cams[0].camera.numpy(), cams[0].eye_pos.numpy()
    rec["R01"], rec["Q01"] = R.numpy(), Q.numpy()
    rec["cam_init_30_16"] = camera.Camera(30, res).camera.numpy()
    np.savez_compressed(os.path.join(HERE, "golden", "shadow.npz"), **rec)
    for f in sorted(os.listdir(os.path.join(HERE, "golden"))):
        print(f, os.path.getsize(os.path.join(HERE, "golden", f)))


if __name__ == "__main__":
    main()
