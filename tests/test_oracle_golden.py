"""Pins oracle/codenerf_oracle.py against fixtures produced by the live reference (oracle/make_golden.py)."""
import os

import numpy as np
import torch

from oracle import codenerf_oracle as orc

T = torch.from_numpy


def load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name), allow_pickle=True)


def close(a, b, tol=1e-6):
    a = a.detach().numpy() if torch.is_tensor(a) else np.asarray(a)
    np.testing.assert_allclose(a, b, rtol=tol, atol=tol)


def test_rays(golden_dir):
    g = load(golden_dir, "rays.npz")
    H, W, n = int(g["H"]), int(g["W"]), int(g["n"])
    d = orc.camera_directions(H, W, T(g["K"]))
    assert np.array_equal(d.numpy(), g["directions"])          # same fp32 ops -> bit exact
    ro, rd = orc.get_bundle(d, T(g["poses"]))
    close(ro, g["ro_full"]); close(rd, g["rd_full"])
    np.random.seed(int(g["np_seed"]))
    sel = orc.draw_select_inds(H * W, 2, n)
    assert np.array_equal(sel, g["select_inds"])               # host RNG stream, bit exact
    ro_s, rd_s = orc.sample_rays(d, T(g["poses"]), sel)
    close(ro_s, g["ro_sel"]); close(rd_s, g["rd_sel"])
    hh, ww = orc.pixel_of_flat_index_np(sel, W)
    assert np.array_equal(rd_s.numpy(), np.concatenate([rd.numpy()[b, hh[b], ww[b]] for b in range(2)]))


def test_point_sampler(golden_dir):
    g = load(golden_dir, "point_sampler.npz")
    Sc, Nf = int(g["Sc"]), int(g["Nf"])
    ro, rd = T(g["ro"]), T(g["rd"])
    for mode in ("lindepth", "lindisp"):
        z, lo, up = orc.depth_tables(Sc, 0.8, 1.8, mode)
        assert np.array_equal(z.numpy(), g[f"z_{mode}"])
        assert np.array_equal(lo.numpy(), g[f"lower_{mode}"]) and np.array_equal(up.numpy(), g[f"upper_{mode}"])
        pts, zv = orc.sample_uniform(ro, rd, z, lo, up, None)
        assert np.array_equal(pts.numpy(), g[f"pts_noperturb_{mode}"])
        assert np.array_equal(zv.numpy(), g[f"zv_noperturb_{mode}"])
    z, lo, up = orc.depth_tables(Sc, 0.8, 1.8, "lindepth")
    pts, zv = orc.sample_uniform(ro, rd, z, lo, up, T(g["t_rand"]))
    assert np.array_equal(pts.numpy(), g["pts_perturb"]) and np.array_equal(zv.numpy(), g["zv_perturb"])
    pts_f, z_f, cdf, inds = orc.sample_pdf(ro, rd, T(g["weights"]), zv, Nf, T(g["u"]))
    assert np.array_equal(z_f.numpy(), g["z_fine"]) and np.array_equal(pts_f.numpy(), g["pts_fine"])
    pts_f, z_f, _, _ = orc.sample_pdf(ro, rd, T(g["weights"]), zv, Nf, None)
    assert np.array_equal(z_f.numpy(), g["z_fine_noperturb"]) and np.array_equal(pts_f.numpy(), g["pts_fine_noperturb"])
    # invariants: sorted, superset of coarse z
    assert np.all(np.diff(z_f.numpy(), axis=-1) >= 0)
    for r in range(z_f.shape[0]):
        assert set(zv.numpy()[r].tolist()) <= set(z_f.numpy()[r].tolist())


def test_searchsorted_indices_bit_exact(golden_dir):
    g = load(golden_dir, "point_sampler.npz")
    assert np.array_equal(orc.searchsorted_right_np(g["ss_cdf"], g["ss_u"]), g["ss_inds"])
    assert g["ss_inds"][0].tolist() == [1, 3, 3, 4, 5, 5]                       # SURVEY section 4 known answer
    # and on the CDFs of the fixture itself
    cdf = orc.pdf_to_cdf(T(g["weights"]))
    want = torch.searchsorted(cdf.contiguous(), T(g["u"]).contiguous(), right=True).numpy()
    assert np.array_equal(orc.searchsorted_right_np(cdf.numpy(), g["u"]), want)


def test_pos_embed(golden_dir):
    g = load(golden_dir, "pos_embed.npz")
    x = T(g["x"])
    for L, logs, inc in ((10, True, True), (4, True, True), (4, False, True), (3, True, False)):
        key = f"{L}_{int(logs)}_{int(inc)}"
        fr = orc.frequency_bands(L, logs)
        assert np.array_equal(fr.numpy(), g["freq_" + key])
        assert np.array_equal(orc.pos_embed(x, fr, inc).numpy(), g["emb_" + key])
    assert orc.frequency_bands(10, True).tolist() == [float(2 ** i) for i in range(10)]


def test_codenerf_mlp_and_forward_pass(golden_dir):
    for tag in ("h64", "h256"):
        g = load(golden_dir, f"codenerf_{tag}.npz")
        hidden, code = int(g["hidden"]), int(g["code"])
        p = {k: v.requires_grad_(True) for k, v in orc.make_codenerf_params(int(g["seed"]), hidden, code).items()}
        zs, zt, x = (T(g[k]).requires_grad_(True) for k in ("z_s", "z_t", "x"))
        y = orc.codenerf_mlp(p, zs, zt, x, 63)
        close(y, g["y"], 2e-5)
        y.backward(T(g["gy"]))
        close(zs.grad, g["g_z_s"], 2e-5); close(zt.grad, g["g_z_t"], 2e-5); close(x.grad, g["g_x"], 2e-5)
        for k, prm in p.items():
            gr = prm.grad.numpy()
            np.testing.assert_allclose(gr.reshape(-1)[::97], g["gprobe_" + k], rtol=1e-4, atol=2e-5)
            np.testing.assert_allclose(gr.astype(np.float64).sum(), float(g["gsum_" + k]), rtol=1e-3, atol=1e-3)
        fx, fd = orc.frequency_bands(10, True), orc.frequency_bands(4, True)
        with torch.no_grad():
            raw = orc.forward_pass(p, T(g["fp_rd"]), T(g["fp_pts"]), T(g["fp_z_s"]), T(g["fp_z_t"]), fx, fd)
            raw_ok = orc.forward_pass(p, T(g["fp_rd"]), T(g["fp_pts"]), T(g["fp_z_s"]), T(g["fp_z_t"]), fx, fd,
                                      pairing="correct")
        close(raw, g["fp_raw"], 2e-5)
        assert not np.allclose(raw_ok.numpy(), g["fp_raw"], atol=1e-4)          # F6: the pairing matters
    assert orc.paired_ray_index_np(5, 3).reshape(-1).tolist() == [0, 1, 2, 3, 4] * 3


def test_volume_render(golden_dir):
    g = load(golden_dir, "volume_render.npz")
    raw, z, rd = (T(g[k]).requires_grad_(True) for k in ("raw", "z", "rd"))
    outs = orc.volume_render(raw, z, rd)
    names = ("rgb", "disp", "acc", "weights", "depth")
    for nme, o in zip(names, outs):
        close(o, g[nme], 1e-6)
    sum((o * T(g["gout_" + nme])).sum() for nme, o in zip(names, outs)).backward()
    close(raw.grad, g["g_raw"], 1e-5); close(z.grad, g["g_z"], 1e-4); close(rd.grad, g["g_rd"], 1e-4)
    acc = outs[2].detach().numpy()
    np.testing.assert_allclose(np.delete(acc, 1), 1.0, atol=1e-6)               # F8: acc == 1 ...
    assert acc[1] < 1e-3                                                         # ... unless softplus(raw-1)*1e10 is tiny (raw=-30 row)


def test_pipeline(golden_dir):
    g = load(golden_dir, "pipeline.npz")
    hidden = int(g["hidden"])
    cfg = orc.RenderConfig(num_coarse=int(g["Sc"]), num_fine=int(g["Nf"]))
    pc = {k: v.requires_grad_(True) for k, v in orc.make_codenerf_params(int(g["seed_c"]), hidden, hidden).items()}
    pf = {k: v.requires_grad_(True) for k, v in orc.make_codenerf_params(int(g["seed_f"]), hidden, hidden).items()}
    ro, rd, zs, zt = (T(g[k]).requires_grad_(True) for k in ("ro", "rd", "z_s", "z_t"))
    rgb_c, rgb_f, aux = orc.predict_radiance_and_render(cfg, ro, rd, pc, pf, zs, zt, T(g["t_rand"]), T(g["u"]))
    close(rgb_c, g["rgb_coarse"], 2e-5); close(rgb_f, g["rgb_fine"], 2e-5)
    tgt = T(g["target"])
    loss = torch.nn.functional.mse_loss(rgb_c, tgt) + torch.nn.functional.mse_loss(rgb_f, tgt)
    np.testing.assert_allclose(loss.item(), float(g["loss"]), rtol=1e-5)
    loss.backward()
    close(ro.grad, g["g_ro"], 2e-5); close(rd.grad, g["g_rd"], 2e-5)
    close(zs.grad, g["g_z_s"], 2e-5); close(zt.grad, g["g_z_t"], 2e-5)
    for tag, p in (("c", pc), ("f", pf)):
        for k, prm in p.items():
            np.testing.assert_allclose(prm.grad.numpy().reshape(-1)[::53], g[f"gprobe_{tag}_{k}"], rtol=1e-4, atol=2e-6)
    assert not aux["z_fine"].requires_grad                                       # F9: sample_pdf has no backward


def test_render_split(golden_dir):
    g = load(golden_dir, "render_split.npz")
    for num_rays, n_gpus, per, pad in g["cases"]:
        p, q = orc.render_split_np(int(num_rays), int(n_gpus))
        assert p.tolist() == list(per) and q.tolist() == list(pad)


def test_flexible_nerf_model(golden_dir):
    """a15: oracle restatement of FlexibleNeRFModel (models/model.py:51-74) and of forward_pass around it, against
    outputs of the live reference class (oracle/make_golden_flex.py)."""
    g = load(golden_dir, "flexnerf.npz")
    for tag in ("small", "lego"):
        L, H, skips = int(g[f"{tag}_L"]), int(g[f"{tag}_H"]), tuple(int(v) for v in g[f"{tag}_skips"])
        p = {k: v.requires_grad_(True) for k, v in orc.make_flexible_nerf_params(31, L, H, skips).items()}
        x = T(g[f"{tag}_x"]).requires_grad_(True)
        y = orc.flexible_nerf_mlp(p, x, 63, L, skips)
        close(y, g[f"{tag}_y"], 2e-5)
        (y * T(g[f"{tag}_gy"])).sum().backward()
        close(x.grad, g[f"{tag}_gx"], 2e-5)
        for name, prm in p.items():
            np.testing.assert_allclose(prm.grad.numpy().reshape(-1)[::97], g[f"{tag}_g_{name}"], rtol=1e-4, atol=2e-5)
        raw = orc.flexible_forward_pass({k: v.detach() for k, v in p.items()}, T(g[f"{tag}_fp_rd"]), T(g[f"{tag}_fp_pts"]),
                                        orc.frequency_bands(10, True), orc.frequency_bands(4, True), L, skips)
        close(raw, g[f"{tag}_fp_raw"], 2e-5)


def test_flexible_nerf_model_without_viewdirs(golden_dir):
    """a15, use_viewdirs=False (fc_out head, models/model.py:47-48,75-76): the direction columns of x carry no gradient."""
    g = load(golden_dir, "flexnerf.npz")
    for tag in ("small_nv", "lego_nv"):
        L, H, skips = int(g[f"{tag}_L"]), int(g[f"{tag}_H"]), tuple(int(v) for v in g[f"{tag}_skips"])
        p = {k: v.requires_grad_(True) for k, v in orc.make_flexible_nerf_params(32, L, H, skips, use_viewdirs=False).items()}
        assert list(p)[-2:] == ["fc_out.weight", "fc_out.bias"] and len(p) == 2 * (L + 1)
        x = T(g[f"{tag}_x"]).requires_grad_(True)
        y = orc.flexible_nerf_mlp(p, x, 63, L, skips)
        close(y, g[f"{tag}_y"], 2e-5)
        (y * T(g[f"{tag}_gy"])).sum().backward()
        close(x.grad, g[f"{tag}_gx"], 2e-5)
        assert float(T(g[f"{tag}_gx"])[:, 63:].abs().max()) == 0.0
        for name, prm in p.items():
            np.testing.assert_allclose(prm.grad.numpy().reshape(-1)[::97], g[f"{tag}_g_{name}"], rtol=1e-4, atol=2e-5)
