"""PSNR delta (BASELINE.json metric, north_star "final PSNR within 0.1 dB"): the same short CodeNeRF training run -- identical
initial weights and codes, identical rays, targets and sample offsets, torch.optim.AdamW with train.py's parameter groups --
once through the drop-in package (libcnrf kernels) and once through the oracle restatement of the reference path executed as
eager torch on the same GPU; then both are evaluated on held-out rays with fixed offsets.

The oracle is the checker here (tests may execute it); it is device-agnostic torch arithmetic, so running it on cuda:0
keeps the comparison to a few seconds at the real layer sizes (hidden 256, codes 256, 32 + 128 samples)."""
import json
import os
import time

import numpy as np
import pytest
import torch

from oracle import codenerf_oracle as orc

pytestmark = pytest.mark.gpu
T = torch.from_numpy

H = W = 64
N_OBJ, RAYS, STEPS, EVAL_RAYS = 4, 1024, 400, 2048
EVAL_AT = (320, 340, 360, 380, 400)      # PSNR is averaged over these checkpoints: a training curve this early is still
                                          # rising ~0.02 dB per step, so a single end point mostly measures trajectory phase
LR = 1e-4                                 # the MLP rate of config/srn-cars-code.yml (codes: 1e-3), as bench.py
SC, NF = 32, 128
TIMES = {}      # seconds per training step (1024 rays incl. H2D of the step's inputs and torch.optim.AdamW), wall clock


def look_at_pose(theta, phi, rho):
    c2w = np.eye(4, dtype=np.float32)
    c2w[0, 0], c2w[1, 0] = -np.sin(phi), np.cos(phi)
    c2w[0, 1], c2w[1, 1], c2w[2, 1] = -np.sin(theta) * np.cos(phi), -np.sin(theta) * np.sin(phi), np.cos(theta)
    c2w[0, 2], c2w[1, 2], c2w[2, 2] = np.cos(theta) * np.cos(phi), np.cos(theta) * np.sin(phi), np.sin(theta)
    c2w[:3, 3] = 1.3 * c2w[:3, 2]
    return c2w


def make_data():
    """Per step: one object, RAYS random pixels of its 64x64 view (shaded ellipse on white), fixed sample offsets."""
    rng = np.random.default_rng(2024)
    K = torch.tensor([[65.6, 0, W / 2, 0], [0, 65.6, H / 2, 0], [0, 0, 1, 0], [0, 0, 0, 1]], dtype=torch.float32)
    dirs = orc.camera_directions(H, W, K)
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    imgs, poses = [], []
    for b in range(N_OBJ):
        cx, cy, r = rng.uniform(24, 40), rng.uniform(24, 40), rng.uniform(12, 18)
        inside = ((xx - cx) ** 2 / r ** 2 + (yy - cy) ** 2 / (0.6 * r) ** 2) < 1.0
        shade = 0.5 + 0.5 * np.cos((xx + yy) / 9.0 + b)
        img = np.stack([np.where(inside, shade * rng.uniform(0.2, 1.0), 1.0) for _ in range(3)], axis=-1).astype(np.float32)
        imgs.append(img.reshape(-1, 3))
        poses.append(look_at_pose(rng.uniform(0.3, 1.2), rng.uniform(-3, 3), 1.3))
    poses = T(np.stack(poses))

    def batch(obj, n):
        sel = rng.permutation(H * W)[:n][None]
        ro, rd = orc.sample_rays(dirs, poses[obj:obj + 1], sel)
        return {"ro": ro.contiguous(), "rd": rd.contiguous(), "tgt": T(imgs[obj][sel[0]]),
                "oid": torch.full((n,), obj, dtype=torch.int64), "t_rand": T(rng.uniform(size=(n, SC)).astype(np.float32)),
                "u": T(rng.uniform(size=(n, NF)).astype(np.float32))}
    train = [batch(s % N_OBJ, RAYS) for s in range(STEPS)]
    held = [batch(o, EVAL_RAYS) for o in range(N_OBJ)]
    return train, held


def init_state():
    g = torch.Generator().manual_seed(11)
    return (orc.make_codenerf_params(1), orc.make_codenerf_params(2), torch.randn(N_OBJ, 256, generator=g),
            torch.randn(N_OBJ, 256, generator=g))


def psnr_of(mse_values):
    return orc.mse2psnr(float(np.mean(mse_values)))


def run_oracle(train, held, dev):
    pc, pf, ts, tt = init_state()
    pc = {k: v.to(dev).requires_grad_(True) for k, v in pc.items()}
    pf = {k: v.to(dev).requires_grad_(True) for k, v in pf.items()}
    ts, tt = ts.to(dev).requires_grad_(True), tt.to(dev).requires_grad_(True)
    cfg = orc.RenderConfig(SC, NF)
    for k in ("freqs_xyz", "freqs_dir", "z", "lower", "upper"):
        setattr(cfg, k, getattr(cfg, k).to(dev))
    opt = torch.optim.AdamW([{"params": list(pc.values())}, {"params": list(pf.values())},
                             {"params": [ts, tt], "lr": 1e-3}], lr=LR)

    def render(b):
        b = {k: v.to(dev) for k, v in b.items()}
        z_s, z_t = orc.embedding_lookup(ts, tt, b["oid"])
        rc, rf, _ = orc.predict_radiance_and_render(cfg, b["ro"], b["rd"], pc, pf, z_s, z_t, b["t_rand"], b["u"])
        return rc, rf, b["tgt"]
    t0, spent, evals = None, 0.0, []
    for i, b in enumerate(train):
        if i == 20:                       # steady state: skip allocator / autotune warm-up
            torch.cuda.synchronize(); t0 = time.perf_counter()
        rc, rf, tgt = render(b)
        loss = torch.nn.functional.mse_loss(rc, tgt) + torch.nn.functional.mse_loss(rf, tgt)
        opt.zero_grad()
        loss.backward()
        opt.step()
        if i + 1 in EVAL_AT:
            torch.cuda.synchronize(); spent -= time.perf_counter()
            with torch.no_grad():
                evals.append(psnr_of([float(torch.nn.functional.mse_loss(render(h)[1], h["tgt"].to(dev))) for h in held]))
            spent += time.perf_counter()
    torch.cuda.synchronize()
    TIMES["reference_path_eager_torch_gpu"] = (time.perf_counter() - t0 - spent) / (len(train) - 20)
    return float(np.mean(evals))


def run_ours(train, held, dev, precision):
    import view_synthesis.nerf as nerf
    import view_synthesis.models as models
    pc, pf, ts, tt = init_state()
    mods = []
    for p in (pc, pf):
        m = models.CodeNeRFModel(256, N_OBJ, 256, 256, 10, 4, True, True)
        m.load_state_dict({k: v.clone() for k, v in p.items()})
        mods.append(m.to(dev))
    emb_tab = models.ShapeTextureEmbedding(N_OBJ, 256, 256)
    emb_tab.shape_embedding.weight.data.copy_(ts)
    emb_tab.texture_embedding.weight.data.copy_(tt)
    emb_tab = emb_tab.to(dev)
    opt = torch.optim.AdamW([{"params": mods[0].parameters()}, {"params": mods[1].parameters()},
                             {"params": emb_tab.parameters(), "lr": 1e-3}], lr=LR)
    ps = nerf.PointSampler(SC, NF, 0.8, 1.8, "lindepth", True, torch.float32, dev)
    emb = (nerf.PositionalEmbedder(10, True, True, torch.float32, dev), nerf.PositionalEmbedder(4, True, True, torch.float32, dev))
    old = nerf.get_precision()
    nerf.set_precision(precision)
    real_rand, real_rand_like = torch.rand, torch.rand_like

    def render(b):
        b = {k: v.to(dev) for k, v in b.items()}
        draws = [b["t_rand"], b["u"]]
        torch.rand = lambda *a, **k: draws.pop(0)              # the samplers' draws (point_sampler.py:64,93), replayed
        torch.rand_like = lambda *a, **k: draws.pop(0)
        try:
            rc, rf = nerf.predict_radiance_and_render((b["ro"], b["rd"]), ps, emb, mods[0], mods[1], emb_tab(b["oid"]))
        finally:
            torch.rand, torch.rand_like = real_rand, real_rand_like
        assert not draws, "a sampler did not consume its injected draw"
        return rc, rf, b["tgt"]
    try:
        t0, spent, evals = None, 0.0, []
        for i, b in enumerate(train):
            if i == 20:
                torch.cuda.synchronize(); t0 = time.perf_counter()
            rc, rf, tgt = render(b)
            loss = torch.nn.functional.mse_loss(rc, tgt) + torch.nn.functional.mse_loss(rf, tgt)
            opt.zero_grad()
            loss.backward()
            opt.step()
            if i + 1 in EVAL_AT:
                torch.cuda.synchronize(); spent -= time.perf_counter()
                with torch.no_grad():
                    evals.append(psnr_of([float(torch.nn.functional.mse_loss(render(h)[1], h["tgt"].to(dev))) for h in held]))
                spent += time.perf_counter()
        torch.cuda.synchronize()
        TIMES[f"ours_{precision}"] = (time.perf_counter() - t0 - spent) / (len(train) - 20)
        return float(np.mean(evals))
    finally:
        nerf.set_precision(old)


def test_psnr_delta_vs_reference_path():
    assert torch.cuda.is_available()
    dev = torch.device("cuda", 0)
    train, held = make_data()
    ref = run_oracle(train, held, dev)
    out = {"steps": STEPS, "rays_per_step": RAYS, "held_out_rays": N_OBJ * EVAL_RAYS, "psnr_reference_path_db": ref}
    for precision in ("fp32", "bf16"):
        ours = run_ours(train, held, dev, precision)
        out[f"psnr_ours_{precision}_db"] = ours
        out[f"delta_{precision}_db"] = ours - ref
    # run-to-run spread of the drop-in itself (weight gradients are accumulated with red.global.add: summation order varies)
    out["psnr_ours_bf16_repeat_db"] = run_ours(train, held, dev, "bf16")
    out["self_spread_bf16_db"] = out["psnr_ours_bf16_repeat_db"] - out["psnr_ours_bf16_db"]
    out["delta_bf16_mean_of_two_runs_db"] = 0.5 * (out["psnr_ours_bf16_db"] + out["psnr_ours_bf16_repeat_db"]) - ref
    out["eval_checkpoints"], out["lr"] = list(EVAL_AT), LR
    out["rays_per_s_training_loop"] = {k: RAYS / v for k, v in TIMES.items()}
    print("PSNR_DELTA " + json.dumps(out))
    path = os.environ.get("CNRF_PSNR_JSON")
    if path:
        json.dump(out, open(path, "w"), indent=1)
    assert ref > 12.0, "the run did not train"                      # white-image baseline of these views is ~8-10 dB
    # north_star: final PSNR within 0.1 dB.  The reference arm is deterministic; the drop-in's own run-to-run spread
    # (self_spread_bf16_db, measured 0.02-0.05 dB) is inside that bound, and the bf16 figure is the mean of two runs.
    assert abs(out["delta_fp32_db"]) <= 0.1 and abs(out["delta_bf16_mean_of_two_runs_db"]) <= 0.1, out
