"""ops.mse_pair: the two photometric losses of train.py:103-104 and their sum in one launch, against
torch.nn.functional.mse_loss -- values and gradients, contiguous and strided (RGBA) operands."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,rgba", [(4096, False), (1000, True), (1, False), (70001, True)])
def test_mse_pair_matches_torch(N, rgba):
    from codenerf_b200 import ops
    assert torch.cuda.is_available()
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(N)
    C = 4 if rgba else 3
    a0 = torch.rand(N, 3, generator=g).to(dev)
    b0 = torch.rand(N, 3, generator=g).to(dev)
    t_full = torch.rand(N, C, generator=g).to(dev)
    tgt = t_full[..., :3]                                    # strided view when rgba
    res = {}
    for mode in ("torch", "ours"):
        a, b = a0.clone().requires_grad_(True), b0.clone().requires_grad_(True)
        if mode == "torch":
            lc = torch.nn.functional.mse_loss(a[..., :3], tgt)
            lf = torch.nn.functional.mse_loss(b[..., :3], tgt)
            loss = lc + lf
        else:
            lc, lf, loss = ops.mse_pair(a, b, tgt)
        (loss + 0.5 * lf).backward()                         # exercises the fine-loss and the sum gradient slots together
        res[mode] = (lc.detach(), lf.detach(), loss.detach(), a.grad, b.grad)
    for x, y in zip(res["torch"], res["ours"]):
        assert x.shape == y.shape
        assert torch.allclose(x, y, rtol=2e-5, atol=1e-9), float((x - y).abs().max())


def test_mse_pair_unused_outputs_and_errors():
    from codenerf_b200 import ops
    dev = torch.device("cuda", 0)
    a = torch.rand(64, 3, device=dev, requires_grad=True)
    b = torch.rand(64, 3, device=dev, requires_grad=True)
    t = torch.rand(64, 3, device=dev)
    lc, _, _ = ops.mse_pair(a, b, t)
    lc.backward()                                            # only the coarse slot has an upstream gradient
    assert torch.allclose(a.grad, 2 * (a.detach() - t) / (64 * 3), rtol=1e-5, atol=1e-9)
    assert float(b.grad.abs().max()) == 0.0
    with pytest.raises(Exception):
        ops.mse_pair(a.cpu(), b.cpu(), t.cpu())              # no CPU fallback


# =====================================================================================================================
# The loss folded into the compositing epilogues (cnrf_render_loss_fwd / _bwd)
# =====================================================================================================================
def _render_setup(dev, N, seed=4):
    import view_synthesis.nerf as nerf
    g = torch.Generator().manual_seed(seed)
    ro = (torch.tensor([0.0, 0.0, 1.3]) + 0.01 * torch.randn(N, 3, generator=g)).to(dev)
    rd = (torch.randn(N, 3, generator=g) * 0.2 + torch.tensor([0.0, 0.0, -1.0])).to(dev)
    t_rand, u = torch.rand(N, 32, generator=g).to(dev), torch.rand(N, 128, generator=g).to(dev)
    ps = nerf.PointSampler(32, 128, 0.8, 1.8, "lindepth", True, torch.float32, dev)
    emb = (nerf.PositionalEmbedder(10, True, True, torch.float32, dev), nerf.PositionalEmbedder(4, True, True, torch.float32, dev))
    return g, ro, rd, t_rand, u, ps, emb


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("target_form", ["rows_rgb", "rows_rgba", "uint8_views"])
@pytest.mark.parametrize("codes", ["table_ids", "broadcast"])
def test_render_loss_equals_render_plus_mse_pair(precision, target_form, codes):
    """predict_radiance_render_and_loss (loss and target gather inside the compositing epilogues) against
    predict_radiance_and_render + ops.mse_pair on the gathered float targets: rgb bit-identical (same kernels), the three
    loss numbers to summation order, every parameter / code gradient to the atomic-accumulation tolerance of the
    fused-vs-staged test.  The uint8 form must reproduce train.py:77-80 + the loader's / 255.0 exactly."""
    import view_synthesis.nerf as nerf
    import view_synthesis.models as models
    dev = torch.device("cuda", 0)
    N = 1500
    g, ro, rd, t_rand, u, ps, emb = _render_setup(dev, N)
    ids = torch.tensor([3, 0, 6]).repeat_interleave(N // 3).to(dev)
    views = torch.randint(0, 256, (5, 24, 24, 4), generator=g, dtype=torch.uint8)
    pixel = torch.randint(0, 5 * 24 * 24, (N,), generator=g, dtype=torch.int64)
    rows_u8 = (views.reshape(-1, 4)[pixel].numpy() / 255.0).astype("float32")          # dataset.py:72,90 + train.py:77-80
    rows = torch.from_numpy(rows_u8).to(dev) if target_form != "rows_rgb" else torch.rand(N, 3, generator=g).to(dev)
    one_code = [torch.randn(1, 256, generator=g), torch.randn(1, 256, generator=g)]
    nerf.set_precision(precision)
    real_rand = torch.rand
    out = {}
    try:
        for mode in ("fused_loss", "staged_loss"):
            torch.manual_seed(21)
            mods = [models.CodeNeRFModel(256, 7, 256, 256, 10, 4).to(dev) for _ in range(2)]
            table = models.ShapeTextureEmbedding(7, 256, 256).to(dev)
            if codes == "table_ids":
                z, leaves = table(ids), list(table.parameters())
            else:
                leaves = [c.clone().to(dev).requires_grad_(True) for c in one_code]
                z = tuple(l.expand(N, -1) for l in leaves)
            q = [t_rand, u]
            torch.rand = lambda *a, **k: q.pop(0)
            if mode == "fused_loss":
                kw = ({"views_u8": views.to(dev), "pixel_index": pixel.to(dev)} if target_form == "uint8_views"
                      else {"target_pixels": rows})
                lc, lf, loss, rc, rf = nerf.predict_radiance_render_and_loss((ro, rd), ps, emb, mods[0], mods[1], z, **kw)
                assert not rc.requires_grad and not rf.requires_grad
            else:
                rc, rf = nerf.predict_radiance_and_render((ro, rd), ps, emb, mods[0], mods[1], z)
                from codenerf_b200 import ops
                lc, lf, loss = ops.mse_pair(rc, rf, rows)
            torch.rand = real_rand
            (loss + 0.5 * lf).backward()
            grads = [p.grad.clone() for m in mods for p in m.parameters()] + [l.grad.clone() for l in leaves]
            out[mode] = (rc.detach().clone(), rf.detach().clone(), torch.stack([lc.detach(), lf.detach(), loss.detach()]), grads)
    finally:
        torch.rand = real_rand
        nerf.set_precision("fp32")
    a, b = out["fused_loss"], out["staged_loss"]
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    assert torch.allclose(a[2], b[2], rtol=2e-5, atol=0), (a[2], b[2])
    ref_fine = torch.nn.functional.mse_loss(b[1], rows[..., :3])
    assert torch.allclose(a[2][1], ref_fine, rtol=2e-5)
    for x, y in zip(a[3], b[3]):
        assert x.shape == y.shape and torch.isfinite(x).all()
        assert torch.allclose(x, y, rtol=1e-4, atol=1e-6 + 1e-4 * float(y.abs().max()))


def test_render_loss_two_calls_and_partial_gradients():
    """Two calls in a row (the accumulators are re-zeroed by every call), a backward through the fine loss only, and the
    pose-gradient fallback (rays that carry gradients take the staged composition)."""
    import view_synthesis.nerf as nerf
    import view_synthesis.models as models
    dev = torch.device("cuda", 0)
    N = 512
    g, ro, rd, t_rand, u, ps, emb = _render_setup(dev, N, seed=9)
    tgt = torch.rand(N, 3, generator=g).to(dev)
    torch.manual_seed(3)
    mods = [models.CodeNeRFModel(256, 2, 256, 256, 10, 4).to(dev) for _ in range(2)]
    z = (torch.randn(N, 256, device=dev), torch.randn(N, 256, device=dev))
    nerf.set_precision("bf16")
    real_rand = torch.rand
    try:
        vals = []
        for _ in range(2):
            q = [t_rand, u]
            torch.rand = lambda *a, **k: q.pop(0)
            lc, lf, loss, rc, rf = nerf.predict_radiance_render_and_loss((ro, rd), ps, emb, mods[0], mods[1], z, target_pixels=tgt)
            torch.rand = real_rand
            vals.append(torch.stack([lc, lf, loss]).detach())
        assert torch.allclose(vals[0], vals[1], rtol=1e-5)
        assert torch.allclose(vals[0][2], torch.nn.functional.mse_loss(rc, tgt) + torch.nn.functional.mse_loss(rf, tgt), rtol=2e-5)
        for m in mods:
            m.zero_grad()
        lf.backward()                                # only the fine term: the coarse model still gets (zero-upstream) gradients
        assert float(mods[1].fc_rgb.weight.grad.abs().max()) > 0
        assert float(mods[0].fc_rgb.weight.grad.abs().max()) == 0.0
        # rays with gradients: staged path, same numbers to bf16-vs-split tolerance is not the point here -- it must run
        ro_g = ro.clone().requires_grad_(True)
        q = [t_rand, u]
        torch.rand = lambda *a, **k: q.pop(0)
        lc2, lf2, loss2, _, _ = nerf.predict_radiance_render_and_loss((ro_g, rd), ps, emb, mods[0], mods[1], z, target_pixels=tgt)
        torch.rand = real_rand
        loss2.backward()
        assert ro_g.grad is not None and torch.isfinite(ro_g.grad).all()
        assert torch.allclose(loss2.detach(), vals[0][2], rtol=5e-2)
    finally:
        torch.rand = real_rand
        nerf.set_precision("fp32")
