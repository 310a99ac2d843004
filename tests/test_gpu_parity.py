"""Parity of the CUDA path (through the drop-in package -> ctypes -> C ABI) against the CPU oracle and the golden
fixtures generated from the live reference.  Run on the B200 box: pytest -m gpu."""
import os

import numpy as np
import pytest
import torch

from oracle import codenerf_oracle as orc

pytestmark = pytest.mark.gpu
T = torch.from_numpy


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda", 0)


def load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name), allow_pickle=True)


def close(a, b, tol=1e-6, rtol=None):
    a = a.detach().cpu().numpy() if torch.is_tensor(a) else np.asarray(a)
    b = b.detach().cpu().numpy() if torch.is_tensor(b) else np.asarray(b)
    np.testing.assert_allclose(a, b, rtol=tol if rtol is None else rtol, atol=tol)


def model_from(params, dev, hidden, code):
    import view_synthesis.models as models
    m = models.CodeNeRFModel(hidden, 1, code, code, 10, 4, True, True)
    m.load_state_dict({k: v.clone() for k, v in params.items()})
    return m.to(dev)


# ---- rays ------------------------------------------------------------------------------------------------
def test_rays_golden(golden_dir, dev):
    import view_synthesis.nerf as nerf
    g = load(golden_dir, "rays.npz")
    H, W, n = int(g["H"]), int(g["W"]), int(g["n"])
    rs = nerf.RaySampler(H, W, T(g["K"]), n, dev, torch.float32)
    assert np.array_equal(rs.directions.cpu().numpy(), g["directions"])            # bit exact
    poses = T(g["poses"]).to(dev)
    ro, rd = rs.get_bundle(poses)
    assert ro.shape == rd.shape == (2, H, W, 3) and ro.stride(1) == 0
    close(ro, g["ro_full"]); close(rd, g["rd_full"])
    np.random.seed(int(g["np_seed"]))
    ro_s, rd_s, sel = rs.sample(poses)
    assert isinstance(sel, np.ndarray) and np.array_equal(sel, g["select_inds"])   # host RNG stream, bit exact
    close(ro_s, g["ro_sel"]); close(rd_s, g["rd_sel"])
    # pixel indexing bit exact: selected rays are exactly rows of the full bundle
    full = rd.reshape(2, H * W, 3)
    assert torch.equal(rd_s, torch.cat([full[b, torch.as_tensor(sel[b]).to(dev)] for b in range(2)]))


def test_rays_backward_to_pose(dev):
    import view_synthesis.nerf as nerf
    H, W, n = 9, 7, 20
    K = torch.tensor([[11.0, 0, 3.3, 0], [0, 11.0, 4.1, 0], [0, 0, 1, 0], [0, 0, 0, 1]])
    rng = np.random.default_rng(0)
    pose_np = rng.normal(size=(3, 4, 4)).astype(np.float32)
    w_ro, w_rd = rng.normal(size=(3 * n, 3)).astype(np.float32), rng.normal(size=(3 * n, 3)).astype(np.float32)
    rs = nerf.RaySampler(H, W, K, n, dev, torch.float32)
    pose = T(pose_np).to(dev).requires_grad_(True)
    np.random.seed(1)
    ro, rd, sel = rs.sample(pose)
    ((ro * T(w_ro).to(dev)).sum() + (rd * T(w_rd).to(dev)).sum()).backward()
    pose_c = T(pose_np).requires_grad_(True)
    ro_c, rd_c = orc.sample_rays(orc.camera_directions(H, W, K), pose_c, sel)
    ((ro_c * T(w_ro)).sum() + (rd_c * T(w_rd)).sum()).backward()
    close(pose.grad, pose_c.grad, 1e-4)
    # full bundle backward
    pose2 = T(pose_np).to(dev).requires_grad_(True)
    ro2, rd2 = rs.get_bundle(pose2)
    (ro2.sum() * 0.5 + (rd2 ** 2).sum()).backward()
    pose3 = T(pose_np).requires_grad_(True)
    ro3, rd3 = orc.get_bundle(orc.camera_directions(H, W, K), pose3)
    (ro3.sum() * 0.5 + (rd3 ** 2).sum()).backward()
    close(pose2.grad, pose3.grad, 1e-3, rtol=1e-4)


# ---- point sampling --------------------------------------------------------------------------------------
def test_point_sampler_golden(golden_dir, dev):
    import view_synthesis.nerf as nerf
    from codenerf_b200 import ops
    g = load(golden_dir, "point_sampler.npz")
    Sc, Nf = int(g["Sc"]), int(g["Nf"])
    ro, rd = T(g["ro"]).to(dev), T(g["rd"]).to(dev)
    for mode in ("lindepth", "lindisp"):
        ps = nerf.PointSampler(Sc, Nf, 0.8, 1.8, mode, False, torch.float32, dev)
        close(ps.z_vals, g[f"z_{mode}"], 1e-7)
        pts, z = ps.sample_uniform(ro, rd)
        close(z, g[f"zv_noperturb_{mode}"], 1e-7); close(pts, g[f"pts_noperturb_{mode}"], 1e-6)
    ps = nerf.PointSampler(Sc, Nf, 0.8, 1.8, "lindepth", True, torch.float32, dev)
    lo, up, zt = (T(g[k]).to(dev) for k in ("lower_lindepth", "upper_lindepth", "z_lindepth"))
    pts, z = ops.sample_uniform(ro, rd, zt, lo, up, T(g["t_rand"]).to(dev))
    assert np.array_equal(z.cpu().numpy(), g["zv_perturb"]) and np.array_equal(pts.cpu().numpy(), g["pts_perturb"])
    # hierarchical, injected u
    w = T(g["weights"]).to(dev)
    pts_f, z_f, cdf, inds = ops.sample_pdf(ro, rd, w, z, Nf, T(g["u"]).to(dev), debug=True)
    close(z_f, g["z_fine"], 2e-6); close(pts_f, g["pts_fine"], 5e-6)
    cdf_ref = orc.pdf_to_cdf(T(g["weights"]))
    close(cdf, cdf_ref, 1e-6)
    pts_f, z_f = ops.sample_pdf(ro, rd, w, z, Nf, None)
    close(z_f, g["z_fine_noperturb"], 2e-6); close(pts_f, g["pts_fine_noperturb"], 5e-6)
    # weights passed as the strided view weights[..., 1:-1] (nerf/__init__.py:87)
    wide = torch.zeros(w.shape[0], Sc, device=dev)
    wide[:, 1:-1] = w
    pts_v, z_v = ops.sample_pdf(ro, rd, wide[..., 1:-1], z, Nf, T(g["u"]).to(dev))
    close(z_v, g["z_fine"], 2e-6)


def test_searchsorted_indices_bit_exact(golden_dir, dev):
    """Bin indices given an identical CDF tensor (north_star): edge table + random CDFs with ties and u on knots."""
    from codenerf_b200 import ops
    g = load(golden_dir, "point_sampler.npz")
    bins = torch.linspace(0, 1, 5)[None].expand(2, -1).contiguous()
    s, inds = ops.invert_cdf(T(g["ss_cdf"]).to(dev), bins.to(dev), T(g["ss_u"]).to(dev))
    assert np.array_equal(inds.cpu().numpy(), g["ss_inds"])
    rng = np.random.default_rng(5)
    N, Cn, K = 257, 63, 128
    w = rng.uniform(0, 1, size=(N, Cn - 1)).astype(np.float32)
    w[rng.uniform(size=w.shape) < 0.3] = 0.0
    cdf = orc.pdf_to_cdf(T(w))
    u = T(rng.uniform(0, 1, size=(N, K)).astype(np.float32))
    u[:, :Cn] = cdf                                          # u exactly on the knots
    zb = torch.sort(T(rng.uniform(0.8, 1.8, size=(N, Cn)).astype(np.float32)), dim=-1)[0]
    want_s, want_i = orc.invert_cdf(cdf, zb, u)
    s, inds = ops.invert_cdf(cdf.to(dev), zb.to(dev), u.to(dev))
    assert torch.equal(inds.cpu(), want_i)                   # bit exact indices
    assert torch.equal(s.cpu(), want_s)                      # and, with unfused arithmetic, bit exact samples
    assert np.array_equal(orc.searchsorted_right_np(cdf.numpy()[:8], u.numpy()[:8]), inds.cpu().numpy()[:8])


@pytest.mark.parametrize("N,Sc,Nf", [(4096, 32, 128), (2048, 64, 64), (1024, 64, 128), (513, 24, 40), (300, 64, 192),
                                     (129, 8, 300)])
def test_sample_pdf_properties_full_size(dev, N, Sc, Nf):
    """C1 shapes (4096 rays, 32+128), C3 (64+64), C5-like (64+128) -- the compile-time-shape kernels -- and odd shapes on the
    runtime-shape kernels (register sort up to 256 new samples, shared-memory sort above): sortedness, coarse depths
    retained, points on the ray; vs oracle values."""
    from codenerf_b200 import ops
    torch.manual_seed(0)
    z_t, lo, up = (t.to(dev) for t in orc.depth_tables(Sc, 0.8, 1.8, "lindepth"))
    ro, rd = torch.randn(N, 3, device=dev), torch.randn(N, 3, device=dev)
    pts_c, z_c = ops.sample_uniform(ro, rd, z_t, lo, up, torch.rand(N, Sc, device=dev))
    w = torch.rand(N, Sc, device=dev) ** 4
    u = torch.rand(N, Nf, device=dev)
    pts, z = ops.sample_pdf(ro, rd, w[..., 1:-1], z_c, Nf, u)
    assert bool((z[:, 1:] >= z[:, :-1]).all())
    merged = torch.sort(torch.cat([z, z_c], dim=-1), dim=-1)[0]
    assert bool((merged[:, 1:] == merged[:, :-1]).sum(dim=-1).ge(Sc).all())      # every coarse depth appears in z
    assert torch.equal(pts, ro[:, None, :] + rd[:, None, :] * z[..., None])
    p_o, z_o, _, _ = orc.sample_pdf(ro.cpu(), rd.cpu(), w[..., 1:-1].cpu(), z_c.cpu(), Nf, u.cpu())
    # the warp-scan CDF differs from torch.cumsum in the last ulp; in bins whose pdf mass is ~1e-5 the lerp
    # divides that ulp by the bin's mass, so a handful of samples move inside their (tiny-probability) bin.
    err = (z.cpu() - z_o).abs()
    assert float((err > 5e-6).float().mean()) < 1e-3 and float(err.max()) < 2e-3


def test_points_backward(dev):
    from codenerf_b200 import ops
    rng = np.random.default_rng(2)
    N, S = 33, 40
    ro_n, rd_n = rng.normal(size=(N, 3)).astype(np.float32), rng.normal(size=(N, 3)).astype(np.float32)
    gp = rng.normal(size=(N, S, 3)).astype(np.float32)
    z_t, lo, up = orc.depth_tables(S, 0.8, 1.8, "lindepth")
    tr = T(rng.uniform(size=(N, S)).astype(np.float32))
    ro, rd = T(ro_n).to(dev).requires_grad_(True), T(rd_n).to(dev).requires_grad_(True)
    pts, z = ops.sample_uniform(ro, rd, z_t.to(dev), lo.to(dev), up.to(dev), tr.to(dev))
    (pts * T(gp).to(dev)).sum().backward()
    ro_c, rd_c = T(ro_n).requires_grad_(True), T(rd_n).requires_grad_(True)
    pts_c, _ = orc.sample_uniform(ro_c, rd_c, z_t, lo, up, tr)
    (pts_c * T(gp)).sum().backward()
    close(ro.grad, ro_c.grad, 1e-5); close(rd.grad, rd_c.grad, 1e-5)


# ---- positional encoding -------------------------------------------------------------------------------------
def test_pos_embed_golden_and_backward(golden_dir, dev):
    import view_synthesis.nerf as nerf
    g = load(golden_dir, "pos_embed.npz")
    for L, logs, inc in ((10, True, True), (4, True, True), (4, False, True), (3, True, False)):
        pe = nerf.PositionalEmbedder(L, logs, inc, torch.float32, dev)
        key = f"{L}_{int(logs)}_{int(inc)}"
        close(pe.frequency_bands, g["freq_" + key], 0)
        x = T(g["x"]).to(dev).requires_grad_(True)
        out = pe.embed(x)
        close(out, g["emb_" + key], 2e-6)
        gout = torch.randn(out.shape, generator=torch.Generator().manual_seed(L))
        (out * gout.to(dev)).sum().backward()
        xc = T(g["x"]).requires_grad_(True)
        (orc.pos_embed(xc, orc.frequency_bands(L, logs), inc) * gout).sum().backward()
        close(x.grad, xc.grad, 1e-3, rtol=1e-5)
    pe = nerf.PositionalEmbedder(4, True, True, torch.float32, dev)
    assert pe.embed(torch.zeros(2, 5, 3, device=dev)).shape == (2, 5, 27)


# ---- compositing -----------------------------------------------------------------------------------------------
def test_volume_render_golden(golden_dir, dev):
    import view_synthesis.nerf as nerf
    g = load(golden_dir, "volume_render.npz")
    raw, z, rd = (T(g[k]).to(dev).requires_grad_(True) for k in ("raw", "z", "rd"))
    outs = nerf.volume_render(raw, z, rd)
    names = ("rgb", "disp", "acc", "weights", "depth")
    for nme, o in zip(names, outs):
        close(o, g[nme], 2e-6, rtol=1e-5)
    sum((o * T(g["gout_" + nme]).to(dev)).sum() for nme, o in zip(names, outs)).backward()
    scale = np.abs(g["g_raw"]).max()
    np.testing.assert_allclose(raw.grad.cpu().numpy(), g["g_raw"], rtol=2e-4, atol=2e-6 * scale)
    np.testing.assert_allclose(z.grad.cpu().numpy(), g["g_z"], rtol=2e-4, atol=2e-6 * np.abs(g["g_z"]).max())
    np.testing.assert_allclose(rd.grad.cpu().numpy(), g["g_rd"], rtol=2e-4, atol=2e-6 * np.abs(g["g_rd"]).max())


@pytest.mark.parametrize("S", [1, 31, 32, 33, 64, 160, 192, 256])
def test_volume_render_vs_oracle_sizes(dev, S):
    import view_synthesis.nerf as nerf
    rng = np.random.default_rng(S)
    N = 50
    raw_n = (rng.normal(size=(N, S, 4)) * 2).astype(np.float32)
    z_n = np.sort(rng.uniform(0.8, 1.8, size=(N, S)).astype(np.float32), axis=-1)
    rd_n = rng.normal(size=(N, 3)).astype(np.float32)
    tgt = T(rng.uniform(size=(N, 3)).astype(np.float32))
    raw, z, rd = (T(a).to(dev).requires_grad_(True) for a in (raw_n, z_n, rd_n))
    rgb, disp, acc, w, depth = nerf.volume_render(raw, z, rd)
    rc, zc, rdc = (T(a).requires_grad_(True) for a in (raw_n, z_n, rd_n))
    rgb_c, disp_c, acc_c, w_c, depth_c = orc.volume_render(rc, zc, rdc)
    close(rgb, rgb_c, 1e-5); close(w, w_c, 1e-5); close(depth, depth_c, 1e-5); close(acc, acc_c, 1e-5)
    close(disp, disp_c, 1e-5, rtol=1e-5)
    (((rgb - tgt.to(dev)) ** 2).mean() + (w[..., 1:-1] ** 2).sum() * 1e-2 if S > 2 else ((rgb - tgt.to(dev)) ** 2).mean()).backward()
    (((rgb_c - tgt) ** 2).mean() + (w_c[..., 1:-1] ** 2).sum() * 1e-2 if S > 2 else ((rgb_c - tgt) ** 2).mean()).backward()
    np.testing.assert_allclose(raw.grad.cpu().numpy(), rc.grad.numpy(), rtol=1e-3, atol=1e-6 * max(1.0, float(rc.grad.abs().max())))
    np.testing.assert_allclose(rd.grad.cpu().numpy(), rdc.grad.numpy(), rtol=1e-3, atol=1e-6 * max(1.0, float(rdc.grad.abs().max())))


def test_volume_render_invariants_full_size(dev):
    """C1 fine-pass shape: weights telescope to 1 (F8) and lie in [0,1]."""
    import view_synthesis.nerf as nerf
    torch.manual_seed(1)
    N, S = 4096, 160
    raw = torch.randn(N, S, 4, device=dev) * 3
    z = torch.sort(torch.rand(N, S, device=dev) + 0.8, dim=-1)[0]
    rd = torch.randn(N, 3, device=dev)
    rgb, disp, acc, w, depth = nerf.volume_render(raw, z, rd)
    close(acc, torch.ones(N), 1e-5)
    close(w.sum(-1), torch.ones(N), 1e-5)
    assert bool((w >= 0).all()) and bool((w <= 1).all())
    assert bool((rgb > -0.0011).all()) and bool((rgb < 1.0011).all())
    assert bool((depth >= z[:, 0] - 1e-5).all()) and bool((depth <= z[:, -1] + 1e-5).all())
