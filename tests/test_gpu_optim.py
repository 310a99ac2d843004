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


@pytest.mark.gpu
def test_fused_adamw_capturable_in_a_cuda_graph():
    """capturable=True: the step count lives on the device (cnrf_adamw_step_dev), so optimiser steps captured in a CUDA graph
    and replayed keep torch.optim.AdamW's arithmetic (bias corrections included) step after step."""
    from codenerf_b200.optim import FusedAdamW
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(1)
    shapes = [(256, 63), (257,), (3, 512), (2458, 256)]
    ref_p = [torch.randn(s, generator=g).to(dev).requires_grad_(True) for s in shapes]
    our_p = [p.detach().clone().requires_grad_(True) for p in ref_p]
    ref = torch.optim.AdamW(ref_p, lr=3e-3, weight_decay=1e-2)
    our = FusedAdamW(our_p, lr=3e-3, weight_decay=1e-2, capturable=True)
    grads = [[torch.randn(s, generator=g).to(dev) for s in shapes] for _ in range(6)]
    static_g = [torch.zeros(s, device=dev) for s in shapes]
    for p, sg in zip(our_p, static_g):
        p.grad = sg
    # two eager steps (state allocation happens here, outside any capture), then four replays of one captured step
    for it in range(2):
        for a, gr, sg in zip(ref_p, grads[it], static_g):
            a.grad = gr.clone(); sg.copy_(gr)
        ref.step(); our.step()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        for sg, gr in zip(static_g, grads[2]):
            sg.copy_(gr)
        snap = [p.detach().clone() for p in our_p]
        states = [(our.state[p]["exp_avg"].clone(), our.state[p]["exp_avg_sq"].clone()) for p in our_p]
        step0 = our._step_dev.clone()
        with torch.cuda.graph(graph, stream=side):
            our.step()
        # capture does not execute: parameters, moments and the device step count are untouched
        assert all(torch.equal(a, b.detach()) for a, b in zip(snap, our_p)) and torch.equal(step0, our._step_dev)
        assert all(torch.equal(s[0], our.state[p]["exp_avg"]) for s, p in zip(states, our_p))
        for it in range(2, 6):
            for a, gr, sg in zip(ref_p, grads[it], static_g):
                a.grad = gr.clone(); sg.copy_(gr)
            ref.step()
            graph.replay()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    for a, b, s in zip(ref_p, our_p, shapes):
        assert torch.allclose(a, b, rtol=2e-6, atol=2e-7), (s, (a - b).abs().max().item())
    assert int(our._step_dev.item()) == 6
    assert float(our.state_dict()["state"][0]["step"]) == 6.0


@pytest.mark.gpu
def test_training_chunk_step_replayed_from_a_cuda_graph():
    """The whole chunk step -- embedding lookup, fused render + loss, backward, FusedAdamW(capturable=True) -- captured once
    with torch.cuda.graph and replayed (bench.py's launch mode) against the same step enqueued from Python: same losses and
    the same parameters after three steps (deterministic sample offsets: perturb off)."""
    import copy
    import view_synthesis.nerf as nerf
    import view_synthesis.models as models
    from codenerf_b200.optim import FusedAdamW
    dev = torch.device("cuda", 0)
    N = 768
    g = torch.Generator().manual_seed(12)
    ro = (torch.tensor([0.0, 0.0, 1.3]) + 0.01 * torch.randn(N, 3, generator=g)).to(dev)
    rd = (torch.randn(N, 3, generator=g) * 0.2 + torch.tensor([0.0, 0.0, -1.0])).to(dev)
    tgt = torch.rand(N, 3, generator=g).to(dev)
    ids = torch.tensor([2, 5, 0]).repeat_interleave(N // 3).to(dev)
    ps = nerf.PointSampler(32, 128, 0.8, 1.8, "lindepth", False, torch.float32, dev)
    emb = (nerf.PositionalEmbedder(10, True, True, torch.float32, dev), nerf.PositionalEmbedder(4, True, True, torch.float32, dev))
    torch.manual_seed(8)
    base = [models.CodeNeRFModel(256, 7, 256, 256, 10, 4).to(dev), models.CodeNeRFModel(256, 7, 256, 256, 10, 4).to(dev),
            models.ShapeTextureEmbedding(7, 256, 256).to(dev)]
    nerf.set_precision("bf16")
    out = {}
    try:
        for mode in ("eager", "graph"):
            mods = [copy.deepcopy(m) for m in base]
            opt = FusedAdamW([{"params": mods[0].parameters()}, {"params": mods[1].parameters()},
                              {"params": mods[2].parameters(), "lr": 1e-3}], lr=1e-4, capturable=(mode == "graph"))
            losses = []

            def step():
                lc, lf, loss, _, _ = nerf.predict_radiance_render_and_loss((ro, rd), ps, emb, mods[0], mods[1], mods[2](ids),
                                                                           target_pixels=tgt)
                opt.zero_grad()
                loss.backward()
                opt.step()
                return loss.detach()

            losses.append(float(step()))                       # eager warm-up step (allocates the optimiser state)
            if mode == "eager":
                for _ in range(3):
                    losses.append(float(step()))
            else:
                torch.cuda.synchronize()
                graph = torch.cuda.CUDAGraph()
                opt.zero_grad(set_to_none=True)
                with torch.cuda.graph(graph):
                    static_loss = step()
                for _ in range(3):
                    graph.replay()
                    losses.append(float(static_loss))
            torch.cuda.synchronize()
            out[mode] = (losses, [p.detach().clone() for m in mods for p in m.parameters()])
    finally:
        nerf.set_precision("fp32")
    (le, pe), (lg, pg) = out["eager"], out["graph"]
    print("\nlosses eager:", le, "graph:", lg)
    assert all(abs(a - b) <= 2e-4 * abs(a) for a, b in zip(le, lg)), (le, lg)
    assert le[-1] < le[0]
    # parameters: AdamW turns the last-bit differences of atomically accumulated gradients into full-size steps wherever a
    # gradient is near zero, so the two runs are compared through their updates (direction and size), not element by element
    p0 = [p.detach() for m in base for p in m.parameters()]
    ue = torch.cat([(a - z).flatten() for a, z in zip(pe, p0)]).double()
    ug = torch.cat([(b - z).flatten() for b, z in zip(pg, p0)]).double()
    cos = float(torch.dot(ue, ug) / (ue.norm() * ug.norm()))
    rel = float((ue - ug).norm() / ue.norm())
    print(f"update after 4 steps, graph vs eager: cosine {cos:.6f}, relative difference {rel:.3e}")
    assert cos > 0.995 and rel < 0.1, (cos, rel)
