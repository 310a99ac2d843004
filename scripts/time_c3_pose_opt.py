"""C3 (config/srn-cars-code-3080-val.yml): eval.py's test-time optimisation loop (eval.py:126-168) -- 2048 rays of a
128 x 128 view per iteration, 64 + 64 samples, codes + camera pose (theta, phi, rho through eval.py's own pose_spherical)
under torch.optim.AdamW -- timed through the drop-in with the pose-gradient chain in its default fp32 parity mode and in the
opt-in bf16 tensor-core mode (CODENERF_POSE_GRAD_PRECISION=bf16), and through the oracle path as eager torch on the same GPU
(BASELINE.md row R2 for this config).  Prints one JSON object.

    python scripts/time_c3_pose_opt.py [iterations=60]
"""
import ast, json, os, sys, time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "code-nerf_b200")); sys.path.insert(0, ROOT)
REF = os.path.join(ROOT, "oracle", "_ref")
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 60
dev = torch.device("cuda", 0)


def reference_pose_spherical():
    path = os.path.join(REF, "eval.py")
    fn = next(n for n in ast.parse(open(path).read()).body if isinstance(n, ast.FunctionDef) and n.name == "pose_spherical")
    ns = {"torch": torch}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), path, "exec"), ns)
    return ns["pose_spherical"]


pose_spherical = reference_pose_spherical()
H = W = 128
n_rays, Sc, Nf, lam = 2048, 64, 64, 1e-6
K = torch.tensor([[131.25, 0, W / 2, 0], [0, 131.25, H / 2, 0], [0, 0, 1, 0], [0, 0, 0, 1]], dtype=torch.float32)
img = torch.rand(H * W, 3, device=dev)


def leaves():
    g = torch.Generator().manual_seed(1)
    return [t.to(dev).requires_grad_(True) for t in (torch.randn(1, 256, generator=g), torch.randn(1, 256, generator=g),
                                                      torch.tensor([1.57]), torch.tensor([0.0]), torch.tensor([1.30]))]


def optimiser(lv):
    return torch.optim.AdamW([{"params": lv[:2]}, {"params": lv[2:4], "lr": 5e-2}, {"params": lv[4:], "lr": 1e-1}], lr=1e-2)


def loss_of(rc, rf, tgt, z_s, z_t):
    return torch.nn.functional.mse_loss(rc[..., :3], tgt) + torch.nn.functional.mse_loss(rf[..., :3], tgt) + \
        lam * (torch.norm(z_s, p=2) + torch.norm(z_t, p=2))


def timed(step, n):
    for _ in range(5):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        step()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n


def drop_in(pose_precision):
    """pose_precision: None (library default), "fp32", "split" or "bf16"."""
    if pose_precision:
        os.environ["CODENERF_POSE_GRAD_PRECISION"] = pose_precision
    else:
        os.environ.pop("CODENERF_POSE_GRAD_PRECISION", None)
    import view_synthesis.nerf as nerf, view_synthesis.models as models
    nerf.set_precision("bf16")
    torch.manual_seed(0); np.random.seed(0)
    mods = [models.CodeNeRFModel(256, 1, 256, 256, 10, 4, True, True).to(dev) for _ in range(2)]
    for m in mods:
        m.requires_grad_(False)                 # eval.py:121-124: the networks are frozen
    rs = nerf.RaySampler(H, W, K, n_rays, dev, torch.float32)
    ps = nerf.PointSampler(Sc, Nf, 0.8, 1.8, "lindepth", True, torch.float32, dev)
    emb = (nerf.PositionalEmbedder(10, True, True, torch.float32, dev), nerf.PositionalEmbedder(4, True, True, torch.float32, dev))
    lv = leaves(); opt = optimiser(lv)

    def step():
        pose = pose_spherical(lv[2], lv[3], lv[4])[None, :]
        ro, rd, sel = rs.sample(tform_cam2world=pose)
        z_s, z_t = lv[0].expand(ro.shape[0], -1), lv[1].expand(ro.shape[0], -1)
        rc, rf = nerf.predict_radiance_and_render((ro, rd), ps, emb, mods[0], mods[1], (z_s, z_t))
        tgt = img[torch.as_tensor(np.asarray(sel).reshape(-1), device=dev)]
        loss = loss_of(rc, rf, tgt, z_s, z_t)
        opt.zero_grad(); loss.backward(); opt.step()
        return loss
    s = timed(step, iters)
    return {"s_per_iteration": s, "rays_per_s": n_rays / s, "final_loss": float(step().detach())}


def eager_reference():
    from oracle import codenerf_oracle as orc
    cfg = orc.RenderConfig(Sc, Nf)
    for k in ("freqs_xyz", "freqs_dir", "z", "lower", "upper"):
        setattr(cfg, k, getattr(cfg, k).to(dev))
    pc = {k: v.to(dev) for k, v in orc.make_codenerf_params(41).items()}
    pf = {k: v.to(dev) for k, v in orc.make_codenerf_params(42).items()}
    dirs = orc.camera_directions(H, W, K).to(dev)
    lv = leaves(); opt = optimiser(lv)

    def step():
        pose = pose_spherical(lv[2], lv[3], lv[4])[None, :]
        sel = np.random.permutation(H * W)[:n_rays][None]
        ro, rd = orc.sample_rays(dirs, pose, sel)
        z_s, z_t = lv[0].expand(n_rays, -1), lv[1].expand(n_rays, -1)
        rc, rf, _ = orc.predict_radiance_and_render(cfg, ro, rd, pc, pf, z_s, z_t, torch.rand(n_rays, Sc, device=dev), torch.rand(n_rays, Nf, device=dev))
        loss = loss_of(rc, rf, img[torch.as_tensor(sel[0], device=dev)], z_s, z_t)
        opt.zero_grad(); loss.backward(); opt.step()
        return loss
    s = timed(step, max(5, iters // 6))
    return {"s_per_iteration": s, "rays_per_s": n_rays / s}


if os.environ.get("C3_ONLY"):          # launch lists: one mode only, e.g. C3_ONLY=split under ncu
    print(json.dumps(drop_in(os.environ["C3_ONLY"]), indent=1))
    sys.exit(0)
out = {"config": "C3: 2048 rays/iteration of a 128x128 view, 64+64 samples, codes + pose (theta, phi, rho) under AdamW, frozen networks",
       "iterations_timed": iters,
       "drop_in_fp32_pose_chain (CODENERF_POSE_GRAD_PRECISION=fp32: CUDA-core FMA)": drop_in("fp32"),
       "drop_in_split_pose_chain (CODENERF_POSE_GRAD_PRECISION=split: split-bf16 tensor-core GEMMs, fp32 activations)": drop_in("split"),
       "drop_in_bf16_pose_chain (CODENERF_POSE_GRAD_PRECISION=bf16: fused chain)": drop_in("bf16"),
       "reference_path_eager_torch_fp32_on_this_gpu": eager_reference()}
print(json.dumps(out, indent=1))
