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
