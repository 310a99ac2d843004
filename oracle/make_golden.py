"""Generate tests/golden/*.npz by running the UNMODIFIED reference from /root/reference on CPU.

Test infrastructure (see oracle/codenerf_oracle.py header).  Run in the build container only
(`python oracle/make_golden.py`): /root/reference does not exist on the GPU box, so the outputs
are committed as small fixtures and this script is committed beside them.

Recipe: stub `imageio` (only the dataset module needs it), put the reference
first on sys.path, import view_synthesis.nerf / view_synthesis.models, drive them with CPU tensors.
Weights come from numpy's PCG64 stream (oracle.make_codenerf_params) so fixtures hold only
inputs and outputs, never 600k-float weight dumps.
"""
import os
import sys
import types

import numpy as np
import torch

REF = os.environ.get("CODENERF_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
OUT = os.path.join(ROOT, "tests", "golden")


def import_reference():
    sys.dont_write_bytecode = True
    sys.modules.setdefault("imageio", types.ModuleType("imageio"))
    for k in [k for k in sys.modules if k == "view_synthesis" or k.startswith("view_synthesis.")]:
        del sys.modules[k]
    sys.path.insert(0, REF)
    import view_synthesis.nerf as nerf
    import view_synthesis.models as models
    sys.path.remove(REF)
    assert os.path.realpath(nerf.__file__).startswith(os.path.realpath(REF)), nerf.__file__
    return nerf, models


def look_at_pose(theta, phi, rho):
    """camera-to-world on a sphere, the closed form of eval.py:22-38 (numpy, no grad)."""
    c2w = np.eye(4, dtype=np.float32)
    c2w[0, 0], c2w[1, 0] = -np.sin(phi), np.cos(phi)
    c2w[0, 1], c2w[1, 1], c2w[2, 1] = -np.sin(theta) * np.cos(phi), -np.sin(theta) * np.sin(phi), np.cos(theta)
    c2w[0, 2], c2w[1, 2], c2w[2, 2] = np.cos(theta) * np.cos(phi), np.cos(theta) * np.sin(phi), np.sin(theta)
    c2w[0, 3], c2w[1, 3], c2w[2, 3] = rho * np.cos(theta) * np.cos(phi), rho * np.cos(theta) * np.sin(phi), rho * np.sin(theta)
    return c2w


def load_params(model, params):
    sd = {k: v.clone() for k, v in params.items()}
    model.load_state_dict(sd)
    return model


def main():
    sys.path.insert(0, ROOT)
    from oracle import codenerf_oracle as orc
    nerf, models = import_reference()
    os.makedirs(OUT, exist_ok=True)
    cpu = torch.device("cpu")
    f32 = torch.float32
    rng = np.random.default_rng(1234)

    # ---- rays -------------------------------------------------------------------------
    H, W, n = 6, 5, 7
    K = torch.tensor([[7.5, 0, 2.25, 0], [0, 7.5, 3.5, 0], [0, 0, 1, 0], [0, 0, 0, 1]], dtype=f32)
    poses = torch.from_numpy(np.stack([look_at_pose(0.6, 0.3, 1.3), look_at_pose(1.1, -2.0, 1.7)]))
    rs = nerf.RaySampler(H, W, K, n, cpu, f32)
    ro_full, rd_full = rs.get_bundle(poses)
    np.random.seed(77)
    ro_s, rd_s, sel = rs.sample(poses)
    np.savez(os.path.join(OUT, "rays.npz"), H=H, W=W, n=n, K=K.numpy(), poses=poses.numpy(),
             directions=rs.directions.numpy(), ro_full=ro_full.contiguous().numpy(), rd_full=rd_full.numpy(),
             np_seed=77, select_inds=sel, ro_sel=ro_s.numpy(), rd_sel=rd_s.numpy())

    # ---- point sampler ----------------------------------------------------------------
    N, Sc, Nf = 9, 8, 16
    ro = torch.from_numpy(rng.normal(size=(N, 3)).astype(np.float32))
    rd = torch.from_numpy(rng.normal(size=(N, 3)).astype(np.float32))
    out = {"N": N, "Sc": Sc, "Nf": Nf, "ro": ro.numpy(), "rd": rd.numpy()}
    for mode in ("lindepth", "lindisp"):
        ps = nerf.PointSampler(Sc, Nf, 0.8, 1.8, mode, False, f32, cpu)
        out[f"z_{mode}"], out[f"lower_{mode}"], out[f"upper_{mode}"] = ps.z_vals.numpy(), ps.lower.numpy(), ps.upper.numpy()
        pts, z = ps.sample_uniform(ro, rd)
        out[f"pts_noperturb_{mode}"], out[f"zv_noperturb_{mode}"] = pts.numpy(), z.contiguous().numpy()
    ps = nerf.PointSampler(Sc, Nf, 0.8, 1.8, "lindepth", True, f32, cpu)
    torch.manual_seed(5)
    pts, z = ps.sample_uniform(ro, rd)
    torch.manual_seed(5)
    out["t_rand"] = torch.rand_like(ps.upper.expand(N, Sc), dtype=f32).numpy()      # replay of point_sampler.py:64
    out["pts_perturb"], out["zv_perturb"] = pts.numpy(), z.numpy()
    # hierarchical: weights with zeros, spikes and ties
    wts = rng.uniform(0, 1, size=(N, Sc - 2)).astype(np.float32)
    wts[0] = 0.0
    wts[1] = 0.0
    wts[1, 2] = 1.0
    wts[2, :3] = 0.0
    wts = torch.from_numpy(wts)
    zc = z
    torch.manual_seed(6)
    pts_f, z_f = ps.sample_pdf(ro, rd, wts, zc)
    torch.manual_seed(6)
    out["u"] = torch.rand([N, Nf], dtype=f32).numpy()                                  # replay of point_sampler.py:93
    out["weights"], out["pts_fine"], out["z_fine"] = wts.numpy(), pts_f.numpy(), z_f.numpy()
    ps_np = nerf.PointSampler(Sc, Nf, 0.8, 1.8, "lindepth", False, f32, cpu)
    pts_f2, z_f2 = ps_np.sample_pdf(ro, rd, wts, zc)
    out["pts_fine_noperturb"], out["z_fine_noperturb"] = pts_f2.numpy(), z_f2.numpy()
    # the searchsorted call site itself (point_sampler.py:101) on the edge table of SURVEY section 4
    cdf = torch.tensor([[0, .2, .2, .7, 1.0], [0, 0, 0.5, 0.5, 1.0]], dtype=f32)
    uu = torch.tensor([[0, .2, .69999, .7, 1, 1.00001], [0, .25, .5, .75, 1.0, -0.1]], dtype=f32)
    out["ss_cdf"], out["ss_u"] = cdf.numpy(), uu.numpy()
    out["ss_inds"] = torch.searchsorted(cdf, uu, right=True).numpy()
    np.savez(os.path.join(OUT, "point_sampler.npz"), **out)

    # ---- positional encoding -------------------------------------------------------------
    x = torch.from_numpy(rng.uniform(-2, 2, size=(11, 3)).astype(np.float32))
    out = {"x": x.numpy()}
    for L, logs, inc in ((10, True, True), (4, True, True), (4, False, True), (3, True, False)):
        pe = nerf.PositionalEmbedder(L, logs, inc, f32, cpu)
        out[f"freq_{L}_{int(logs)}_{int(inc)}"] = pe.frequency_bands.numpy()
        out[f"emb_{L}_{int(logs)}_{int(inc)}"] = pe.embed(x).numpy()
    np.savez(os.path.join(OUT, "pos_embed.npz"), **out)

    # ---- CodeNeRFModel fwd/bwd + forward_pass ---------------------------------------------
    for tag, hidden, code in (("h64", 64, 64), ("h256", 256, 256)):
        p = orc.make_codenerf_params(seed=11, hidden=hidden, code=code)
        model = load_params(models.CodeNeRFModel(hidden, 1, code, code, 10, 4, True, True), p)
        M = 10
        zs = torch.from_numpy(rng.normal(size=(M, code)).astype(np.float32)).requires_grad_(True)
        zt = torch.from_numpy(rng.normal(size=(M, code)).astype(np.float32)).requires_grad_(True)
        xin = torch.from_numpy(rng.uniform(-1, 1, size=(M, 90)).astype(np.float32)).requires_grad_(True)
        y = model(zs, zt, xin)
        gy = torch.from_numpy(rng.normal(size=(M, 4)).astype(np.float32))
        y.backward(gy)
        out = {"hidden": hidden, "code": code, "seed": 11, "z_s": zs.detach().numpy(), "z_t": zt.detach().numpy(),
               "x": xin.detach().numpy(), "y": y.detach().numpy(), "gy": gy.numpy(), "g_z_s": zs.grad.numpy(),
               "g_z_t": zt.grad.numpy(), "g_x": xin.grad.numpy()}
        for name, prm in model.named_parameters():
            g = prm.grad.numpy()
            # fixtures stay small: keep a strided probe of every weight grad plus its full sum
            out["gsum_" + name] = np.float64(g.astype(np.float64).sum())
            out["gprobe_" + name] = g.reshape(-1)[::97].copy()
        # forward_pass with the F6 pairing
        Nr, S = 5, 3
        pe_x, pe_d = nerf.PositionalEmbedder(10, True, True, f32, cpu), nerf.PositionalEmbedder(4, True, True, f32, cpu)
        rdd = torch.from_numpy(rng.normal(size=(Nr, 3)).astype(np.float32))
        pts = torch.from_numpy(rng.uniform(-1, 1, size=(Nr, S, 3)).astype(np.float32))
        zs2 = torch.from_numpy(rng.normal(size=(Nr, code)).astype(np.float32))
        zt2 = torch.from_numpy(rng.normal(size=(Nr, code)).astype(np.float32))
        with torch.no_grad():
            raw = nerf.forward_pass(model, (pe_x, pe_d), rdd, pts, (zs2, zt2))
        out.update({"fp_rd": rdd.numpy(), "fp_pts": pts.numpy(), "fp_z_s": zs2.numpy(), "fp_z_t": zt2.numpy(),
                    "fp_raw": raw.numpy()})
        np.savez(os.path.join(OUT, f"codenerf_{tag}.npz"), **out)

    # ---- volume_render fwd/bwd ---------------------------------------------------------------
    N, S = 7, 12
    raw = torch.from_numpy((rng.normal(size=(N, S, 4)) * 3).astype(np.float32)).requires_grad_(True)
    raw.data[0, :, 3] = 30.0      # softplus linear branch (threshold 20)
    raw.data[1, :, 3] = -30.0     # nearly empty ray
    z = torch.sort(torch.from_numpy(rng.uniform(0.8, 1.8, size=(N, S)).astype(np.float32)), dim=-1)[0].requires_grad_(True)
    rd = torch.from_numpy(rng.normal(size=(N, 3)).astype(np.float32)).requires_grad_(True)
    outs = nerf.volume_render(raw, z, rd)
    gs = [torch.from_numpy(rng.normal(size=tuple(o.shape)).astype(np.float32)) for o in outs]
    loss = sum((o * g).sum() for o, g in zip(outs, gs))
    loss.backward()
    names = ("rgb", "disp", "acc", "weights", "depth")
    out = {"raw": raw.detach().numpy(), "z": z.detach().numpy(), "rd": rd.detach().numpy(),
           "g_raw": raw.grad.numpy(), "g_z": z.grad.numpy(), "g_rd": rd.grad.numpy()}
    for nme, o, g in zip(names, outs, gs):
        out[nme], out["gout_" + nme] = o.detach().numpy(), g.numpy()
    np.savez(os.path.join(OUT, "volume_render.npz"), **out)

    # ---- predict_radiance_and_render end to end (perturb on, RNG replayed) -------------------
    Nr, Sc, Nf, hidden = 12, 8, 16, 64
    pc, pf = orc.make_codenerf_params(21, hidden, hidden), orc.make_codenerf_params(22, hidden, hidden)
    mc = load_params(models.CodeNeRFModel(hidden, 1, hidden, hidden, 10, 4, True, True), pc)
    mf = load_params(models.CodeNeRFModel(hidden, 1, hidden, hidden, 10, 4, True, True), pf)
    ps = nerf.PointSampler(Sc, Nf, 0.8, 1.8, "lindepth", True, f32, cpu)
    pe_x, pe_d = nerf.PositionalEmbedder(10, True, True, f32, cpu), nerf.PositionalEmbedder(4, True, True, f32, cpu)
    pose = torch.from_numpy(look_at_pose(0.9, 0.4, 1.3))[None]
    rs = nerf.RaySampler(8, 8, torch.tensor([[8.2, 0, 4, 0], [0, 8.2, 4, 0], [0, 0, 1, 0], [0, 0, 0, 1]], dtype=f32), Nr, cpu, f32)
    np.random.seed(3)
    ro, rd, sel = rs.sample(pose)
    ro, rd = ro.clone().requires_grad_(True), rd.clone().requires_grad_(True)
    zs = torch.from_numpy(rng.normal(size=(Nr, hidden)).astype(np.float32)).requires_grad_(True)
    zt = torch.from_numpy(rng.normal(size=(Nr, hidden)).astype(np.float32)).requires_grad_(True)
    tgt = torch.from_numpy(rng.uniform(0, 1, size=(Nr, 3)).astype(np.float32))
    torch.manual_seed(9)
    rgb_c, rgb_f = nerf.predict_radiance_and_render((ro, rd), ps, (pe_x, pe_d), mc, mf, (zs, zt))
    torch.manual_seed(9)
    t_rand = torch.rand_like(ps.upper.expand(Nr, Sc), dtype=f32)
    u = torch.rand([Nr, Nf], dtype=f32)
    loss = torch.nn.functional.mse_loss(rgb_c, tgt) + torch.nn.functional.mse_loss(rgb_f, tgt)
    loss.backward()
    out = {"Nr": Nr, "Sc": Sc, "Nf": Nf, "hidden": hidden, "seed_c": 21, "seed_f": 22, "ro": ro.detach().numpy(),
           "rd": rd.detach().numpy(), "z_s": zs.detach().numpy(), "z_t": zt.detach().numpy(), "target": tgt.numpy(),
           "t_rand": t_rand.numpy(), "u": u.numpy(), "rgb_coarse": rgb_c.detach().numpy(),
           "rgb_fine": rgb_f.detach().numpy(), "loss": np.float64(loss.item()), "g_ro": ro.grad.numpy(),
           "g_rd": rd.grad.numpy(), "g_z_s": zs.grad.numpy(), "g_z_t": zt.grad.numpy()}
    for tag, m in (("c", mc), ("f", mf)):
        for name, prm in m.named_parameters():
            g = prm.grad.numpy()
            out[f"gsum_{tag}_{name}"] = np.float64(g.astype(np.float64).sum())
            out[f"gprobe_{tag}_{name}"] = g.reshape(-1)[::53].copy()
    np.savez(os.path.join(OUT, "pipeline.npz"), **out)

    # ---- parallel_image_render split arithmetic (nerf/__init__.py:179-187) ----------------
    rows = []
    for num_rays, n_gpus in ((16384, 8), (9216, 7), (9216, 8), (640000, 8), (100, 3), (7, 2), (5, 1)):
        per = torch.full([n_gpus], (num_rays / n_gpus), dtype=int)
        padding = num_rays - torch.sum(per)
        per[-1] = num_rays - torch.sum(per[:-1])
        pad = torch.zeros([n_gpus], dtype=int)
        if padding > 0:
            pad[:-1] = padding
        rows.append((num_rays, n_gpus, per.numpy().tolist(), pad.numpy().tolist()))
    np.savez(os.path.join(OUT, "render_split.npz"), cases=np.array(rows, dtype=object), allow_pickle=True)
    print("wrote", sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
