"""FusedAdamW (csrc/adamw.cu) against torch.optim.AdamW on the same parameters and gradients."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "code-nerf_b200"))


@pytest.mark.gpu
def test_fused_adamw_matches_torch_adamw():
    from codenerf_b200.optim import FusedAdamW
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(0)
    shapes = [(256, 63), (256,), (257, 512), (3, 512), (3,), (2458, 256), (1,), (5, 7)]       # ragged, unaligned tails
    ref_p = [torch.randn(s, generator=g).to(dev).requires_grad_(True) for s in shapes]
    our_p = [p.detach().clone().requires_grad_(True) for p in ref_p]
    groups = lambda ps: [{"params": ps[:5]}, {"params": ps[5:], "lr": 1e-3, "weight_decay": 0.0}]
    ref = torch.optim.AdamW(groups(ref_p), lr=1e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2)
    our = FusedAdamW(groups(our_p), lr=1e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2)
    for it in range(5):
        for a, b in zip(ref_p, our_p):
            if it == 2 and a.dim() == 1:
                a.grad = b.grad = None                      # a parameter without a gradient skips its step count
                continue
            gr = torch.randn(a.shape, generator=g).to(dev) * (10.0 ** (it - 2))
            a.grad, b.grad = gr.clone(), gr.clone()
        ref.step(); our.step()
    for a, b, s in zip(ref_p, our_p, shapes):
        assert torch.allclose(a, b, rtol=2e-6, atol=2e-7), (s, (a - b).abs().max().item())
    sd = our.state_dict()
    assert set(sd["state"][0].keys()) == {"step", "exp_avg", "exp_avg_sq"}
    ref2 = torch.optim.AdamW(groups([p.detach().clone().requires_grad_(True) for p in our_p]), lr=1e-4)
    ref2.load_state_dict(sd)                                 # checkpoints interchange with torch.optim.AdamW
    assert float(ref2.state_dict()["state"][0]["step"]) == float(sd["state"][0]["step"])


@pytest.mark.gpu
def test_fused_adamw_refuses_cpu_parameters():
    from codenerf_b200.optim import FusedAdamW
    from codenerf_b200 import _lib
    p = torch.zeros(4, requires_grad=True)
    p.grad = torch.ones(4)
    with pytest.raises(_lib.CnrfError):
        FusedAdamW([p]).step()


@pytest.mark.gpu
def test_fused_adamw_more_than_one_batch_with_an_empty_tensor():
    """More parameters than one launch batch holds (40) with a zero-element tensor among them: every tensor must be stepped
    exactly once (the batch loop resumes at the first row the previous batch did not consume)."""
    from codenerf_b200.optim import FusedAdamW
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    shapes = [(7,), (0,), (33, 5)] + [(3 + i,) for i in range(50)]
    ps = [torch.randn(*s, device=dev).requires_grad_(True) for s in shapes]
    ref = [p.detach().clone().requires_grad_(True) for p in ps]
    a, b = FusedAdamW(ps, lr=1e-2), torch.optim.AdamW(ref, lr=1e-2)
    for _ in range(3):
        for p, r in zip(ps, ref):
            g = torch.randn_like(p)
            p.grad, r.grad = g.clone(), g.clone()
        a.step(); b.step()
    for p, r in zip(ps, ref):
        assert torch.allclose(p, r, rtol=1e-5, atol=1e-6)
