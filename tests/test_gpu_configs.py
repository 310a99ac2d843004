"""BASELINE.json configs[2] and [3] as parity cases (configs[1] is the bench, [4] is tests/test_gpu_flexnerf.py):

* C3 srn-cars-code-3080-val -- eval.py:120-168: test-time optimisation of one broadcast (stride-0) code pair AND the camera
  pose (theta, phi, rho through pose_spherical -> RaySampler.sample -> the whole render path -> two MSE losses + the code-norm
  regulariser), 64 + 64 samples; the loss trajectory and the first-step gradients of all five leaves against the CPU oracle.
* C4 srn-chairs-code -- near/far 1.25/2.75 and chairs-sized / stress-sized embedding tables through the shared-code path:
  dense table gradients equal what torch.nn.Embedding's backward accumulates for the per-ray path.
"""
import numpy as np
import pytest
import torch

from oracle import codenerf_oracle as orc

pytestmark = pytest.mark.gpu
T = torch.from_numpy


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda", 0)


def pose_spherical(theta, phi, rho):
    """eval.py:22-38, written without in-place assignment (identical values and gradients)."""
    z = torch.zeros_like(theta)
    o = torch.ones_like(theta)
    rows = [torch.cat([-torch.sin(phi), -torch.sin(theta) * torch.cos(phi), torch.cos(theta) * torch.cos(phi),
                       rho * torch.cos(theta) * torch.cos(phi)]),
            torch.cat([torch.cos(phi), -torch.sin(theta) * torch.sin(phi), torch.cos(theta) * torch.sin(phi),
                       rho * torch.cos(theta) * torch.sin(phi)]),
            torch.cat([z, torch.cos(theta), torch.sin(theta), rho * torch.sin(theta)]),
            torch.cat([z, z, z, o])]
    return torch.stack(rows)


def test_c3_test_time_code_and_pose_optimisation(dev):
    import view_synthesis.nerf as nerf
    import view_synthesis.models as models
    Hh = Ww = 32
    n_rays, Sc, Nf, steps = 192, 64, 64, 4
    rng = np.random.default_rng(33)
    K = torch.tensor([[32.8, 0, Ww / 2, 0], [0, 32.8, Hh / 2, 0], [0, 0, 1, 0], [0, 0, 0, 1]], dtype=torch.float32)
    pc, pf = orc.make_codenerf_params(41), orc.make_codenerf_params(42)
    img = T(rng.uniform(0, 1, size=(Hh * Ww, 3)).astype(np.float32))
    sels = [rng.permutation(Hh * Ww)[:n_rays][None] for _ in range(steps)]
    draws = [(T(rng.uniform(size=(n_rays, Sc)).astype(np.float32)), T(rng.uniform(size=(n_rays, Nf)).astype(np.float32)))
             for _ in range(steps)]
    zs0, zt0 = T(rng.normal(size=(1, 256)).astype(np.float32)), T(rng.normal(size=(1, 256)).astype(np.float32))
    lam = 1e-4

    def leaves(device):
        return [t.clone().to(device).requires_grad_(True) for t in
                (zs0, zt0, torch.tensor([1.2]), torch.tensor([0.3]), torch.tensor([1.3]))]

    def optimiser(lv):       # eval.py:133-138 (AdamW, separate angle / radius rates)
        return torch.optim.AdamW([{"params": lv[:2]}, {"params": lv[2:4], "lr": 1e-2}, {"params": lv[4:], "lr": 1e-2}], lr=1e-2)

    # ---- CPU oracle
    cfg = orc.RenderConfig(Sc, Nf)
    dirs = orc.camera_directions(Hh, Ww, K)
    lv_o = leaves("cpu")
    opt_o = optimiser(lv_o)
    losses_o, grads_o = [], None
    for it in range(steps):
        pose = pose_spherical(lv_o[2], lv_o[3], lv_o[4])[None]
        ro, rd = orc.sample_rays(dirs, pose, sels[it])
        z_s, z_t = lv_o[0].expand(n_rays, -1), lv_o[1].expand(n_rays, -1)
        rc, rf, _ = orc.predict_radiance_and_render(cfg, ro, rd, pc, pf, z_s, z_t, *draws[it])
        tgt = img[sels[it][0]]
        loss = torch.nn.functional.mse_loss(rc, tgt) + torch.nn.functional.mse_loss(rf, tgt) + \
            lam * (torch.norm(z_s, p=2) + torch.norm(z_t, p=2))
        opt_o.zero_grad()
        loss.backward()
        if it == 0:
            grads_o = [l.grad.clone() for l in lv_o]
        opt_o.step()
        losses_o.append(float(loss.detach()))

    # ---- drop-in on the GPU (default precision: the pose gradient request selects the fp32 chain by itself)
    mods = []
    for p in (pc, pf):
        m = models.CodeNeRFModel(256, 1, 256, 256, 10, 4, True, True)
        m.load_state_dict({k: v.clone() for k, v in p.items()})
        mods.append(m.to(dev))
    rs = nerf.RaySampler(Hh, Ww, K, n_rays, dev, torch.float32)
    ps = nerf.PointSampler(Sc, Nf, 0.8, 1.8, "lindepth", True, torch.float32, dev)
    emb = (nerf.PositionalEmbedder(10, True, True, torch.float32, dev), nerf.PositionalEmbedder(4, True, True, torch.float32, dev))
    lv = leaves(dev)
    opt = optimiser(lv)
    losses, grads = [], None
    real_rand, real_perm = torch.rand, np.random.permutation
    try:
        for it in range(steps):
            pose = pose_spherical(lv[2], lv[3], lv[4])[None]
            np.random.permutation = lambda a, _s=sels[it]: np.concatenate([_s[0], np.setdiff1d(np.arange(np.size(a) if np.ndim(a) else a), _s[0])])
            ro, rd, sel = rs.sample(tform_cam2world=pose)                 # eval.py:146
            assert np.array_equal(np.asarray(sel).reshape(-1), sels[it][0])
            z_s, z_t = lv[0].expand(ro.shape[0], -1), lv[1].expand(ro.shape[0], -1)          # stride-0 codes, eval.py:149
            q = [d.to(dev) for d in draws[it]]
            torch.rand = lambda *a, **k: q.pop(0)
            rc, rf = nerf.predict_radiance_and_render((ro, rd), ps, emb, mods[0], mods[1], (z_s, z_t))
            torch.rand = real_rand
            tgt = img[sels[it][0]].to(dev)
            loss = torch.nn.functional.mse_loss(rc[..., :3], tgt) + torch.nn.functional.mse_loss(rf[..., :3], tgt) + \
                lam * (torch.norm(z_s, p=2) + torch.norm(z_t, p=2))
            opt.zero_grad()
            loss.backward()
            if it == 0:
                grads = [l.grad.clone().cpu() for l in lv]
            opt.step()
            losses.append(float(loss.detach()))
    finally:
        torch.rand, np.random.permutation = real_rand, real_perm
    np.testing.assert_allclose(losses[0], losses_o[0], rtol=1e-4)
    for got, ref, name in zip(grads, grads_o, ("z_s", "z_t", "theta", "phi", "rho")):
        np.testing.assert_allclose(got.numpy(), ref.numpy(), rtol=5e-3, atol=2e-5 * max(1.0, float(ref.abs().max())), err_msg=name)
    np.testing.assert_allclose(losses, losses_o, rtol=2e-3)               # four coupled AdamW steps stay together


@pytest.mark.parametrize("num_objects", [4612, 50000])
def test_c4_chairs_tables_shared_codes(dev, num_objects):
    """srn-chairs-code: near/far 1.25/2.75, large embedding tables; four images per chunk (runs of equal ids)."""
    import view_synthesis.nerf as nerf
    import view_synthesis.models as models
    torch.manual_seed(5)
    N, Sc, Nf = 1024, 32, 128
    mods = [models.CodeNeRFModel(256, num_objects, 256, 256, 10, 4).to(dev) for _ in range(2)]
    table = models.ShapeTextureEmbedding(num_objects, 256, 256).to(dev)
    ps = nerf.PointSampler(Sc, Nf, 1.25, 2.75, "lindepth", True, torch.float32, dev)
    emb = (nerf.PositionalEmbedder(10, True, True, torch.float32, dev), nerf.PositionalEmbedder(4, True, True, torch.float32, dev))
    ids = torch.tensor([num_objects - 1, 0, 1234, num_objects // 2], device=dev).repeat_interleave(N // 4)
    ro = torch.tensor([0.0, 0.0, 2.0], device=dev).expand(N, 3).contiguous()
    rd = torch.nn.functional.normalize(torch.randn(N, 3, device=dev) * 0.2 + torch.tensor([0.0, 0.0, -1.0], device=dev), dim=-1)
    tgt = torch.rand(N, 3, device=dev)
    t_rand, u = torch.rand(N, Sc, device=dev), torch.rand(N, Nf, device=dev)
    old = nerf.get_precision()
    nerf.set_precision("bf16")
    real_rand = torch.rand
    out = {}
    try:
        for mode in ("shared", "per_ray"):
            for m in mods + [table]:
                m.zero_grad()
            codes = table(ids)
            if mode == "per_ray":       # plain tensors: no code-source tag, torch.nn.Embedding's own backward
                codes = tuple(c * 1.0 for c in codes)
            q = [t_rand, u]
            torch.rand = lambda *a, **k: q.pop(0)
            rc, rf = nerf.predict_radiance_and_render((ro, rd), ps, emb, mods[0], mods[1], codes)
            torch.rand = real_rand
            (torch.nn.functional.mse_loss(rc, tgt) + torch.nn.functional.mse_loss(rf, tgt)).backward()
            out[mode] = (rf.detach().clone(), table.shape_embedding.weight.grad.clone(), table.texture_embedding.weight.grad.clone())
    finally:
        torch.rand = real_rand
        nerf.set_precision(old)
    a, b = out["shared"], out["per_ray"]
    assert torch.allclose(a[0], b[0], atol=2e-3)
    used = torch.unique(ids)
    for ga, gb in ((a[1], b[1]), (a[2], b[2])):
        assert ga.shape == (num_objects, 256)
        mask = torch.ones(num_objects, dtype=torch.bool, device=dev)
        mask[used] = False
        assert float(ga[mask].abs().max()) == 0.0 and float(gb[mask].abs().max()) == 0.0      # untouched rows stay zero
        scale = float(gb[used].abs().max())
        assert scale > 0
        assert float((ga[used] - gb[used]).abs().max()) <= 3e-2 * scale
        cos = torch.nn.functional.cosine_similarity(ga[used].flatten(), gb[used].flatten(), dim=0)
        assert float(cos) > 0.995
