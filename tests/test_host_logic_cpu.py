"""Host-side logic of the drop-in package (no GPU): module surface, parameter names, split arithmetic, gloo gather."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import codenerf_oracle as orc


def test_module_surface_matches_reference_names():
    import view_synthesis.nerf as nerf
    import view_synthesis.models as models
    for n in ("RaySampler", "PointSampler", "PositionalEmbedder", "volume_render", "prepare_samplers",
              "prepare_embedders", "predict_radiance_and_render", "forward_pass", "parallel_image_render"):
        assert hasattr(nerf, n)
    for n in ("CodeNeRFModel", "ShapeTextureEmbedding", "get_params_tensor"):
        assert hasattr(models, n)
    import view_synthesis
    assert os.path.realpath(nerf.__file__).startswith(os.path.realpath(os.path.join(os.path.dirname(__file__), "..", "code-nerf_b200")))


def test_state_dict_names_and_shapes():
    import view_synthesis.models as models
    m = models.CodeNeRFModel(256, 1, 256, 256, 10, 4, True, True)
    sd = m.state_dict()
    shapes = orc.codenerf_shapes(256, 256, 63, 27)
    assert sorted(sd) == sorted([f"{k}.weight" for k in shapes] + [f"{k}.bias" for k in shapes])
    for k, (o, i) in shapes.items():
        assert tuple(sd[k + ".weight"].shape) == (o, i) and tuple(sd[k + ".bias"].shape) == (o,)
    assert sum(p.numel() for p in m.parameters()) == 616964
    e = models.ShapeTextureEmbedding(7, 256, 256)
    assert sorted(e.state_dict()) == ["shape_embedding.weight", "texture_embedding.weight"]
    s, t = e.get_params_tensor()
    assert s.numel() == 7 * 256 and not s.requires_grad                          # F10: detached regulariser
    assert models.get_params_tensor(e, False)[1].numel() == 7 * 256


def test_constructor_asserts():
    import view_synthesis.nerf as nerf
    cpu, f32 = torch.device("cpu"), torch.float32
    with pytest.raises(AssertionError):
        nerf.PointSampler(8, 8, 2.0, 1.0, "lindepth", False, f32, cpu)
    with pytest.raises(AssertionError):
        nerf.PositionalEmbedder(0, True, True, f32, cpu)
    with pytest.raises(AssertionError):
        nerf.RaySampler(4, 4, torch.eye(4), 17, cpu, f32)
    with pytest.raises(AssertionError):
        nerf.RaySampler(4, 4, torch.eye(3), 4, cpu, f32)
    ps = nerf.PointSampler(8, 16, 0.8, 1.8, "lindepth", True, f32, cpu)
    z, lo, up = orc.depth_tables(8, 0.8, 1.8, "lindepth")
    assert torch.equal(ps.z_vals, z) and torch.equal(ps.lower, lo) and torch.equal(ps.upper, up)
    pe = nerf.PositionalEmbedder(10, True, True, f32, cpu)
    assert pe.frequency_bands.tolist() == [float(2 ** i) for i in range(10)]


def test_rng_draw_is_the_reference_draw():
    """sample_uniform draws torch.rand((N,S)); the reference draws rand_like(upper.expand(N,S)) (point_sampler.py:64)."""
    up = torch.arange(5.0)
    torch.manual_seed(3)
    a = torch.rand_like(up.expand(7, 5), dtype=torch.float32)
    torch.manual_seed(3)
    b = torch.rand((7, 5), dtype=torch.float32)
    assert torch.equal(a, b)


def test_render_split(golden_dir):
    import view_synthesis.nerf as nerf
    g = np.load(os.path.join(golden_dir, "render_split.npz"), allow_pickle=True)
    for num_rays, n_gpus, per, pad in g["cases"]:
        p, q = nerf.render_split(int(num_rays), int(n_gpus))
        assert p == list(per) and q == list(pad)


class _Cfg(dict):
    __getattr__ = dict.__getitem__


def _fake_render(rays, point_sampler, embedders, mc, mf, latent):
    ro, rd = rays
    return None, rd * 2.0 + ro


def _worker(rank, world, port, H, W, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import view_synthesis.nerf as nerf
    nerf.predict_radiance_and_render = _fake_render

    class FakeRays:
        def get_bundle(self, tform_cam2world):
            rd = torch.arange(H * W * 3, dtype=torch.float32).reshape(1, H, W, 3)
            return torch.ones_like(rd), rd
    cfg = _Cfg(is_distributed=True, gpus=world, nerf=_Cfg(validation=_Cfg(chunksize=5)))
    models = {"nerf_coarse": torch.nn.Identity(), "nerf_fine": torch.nn.Identity()}
    rgb = nerf.parallel_image_render(cfg, torch.eye(4)[None], [torch.zeros(1, 4), torch.zeros(1, 4)], models,
                                     (FakeRays(), None), (None, None), torch.device("cpu"))
    if rank == 0:
        torch.save(rgb, out)
    else:
        assert rgb is None
    dist.destroy_process_group()


@pytest.mark.parametrize("H,W", [(3, 5), (4, 4)])
def test_parallel_image_render_two_ranks_gloo(tmp_path, H, W):
    out = str(tmp_path / "rgb.pt")
    port = 29500 + (os.getpid() % 2000) + H
    mp.spawn(_worker, args=(2, port, H, W, out), nprocs=2, join=True)
    rgb = torch.load(out)
    want = torch.arange(H * W * 3, dtype=torch.float32).reshape(-1, 3) * 2.0 + 1.0
    assert torch.equal(rgb, want)          # ordering and trimming of the padded all_gather (nerf/__init__.py:212-226)


def test_embedding_output_carries_its_source_and_equals_nn_embedding():
    """ShapeTextureEmbedding.forward returns what the reference returns (two nn.Embedding lookups, models/model.py:100-103)
    and tags both tensors with (ids, shape table, texture table) so the fused field can share the code layers between
    rays of one object; slicing or any other op drops the tag (-> per-ray path)."""
    import view_synthesis.models as models
    torch.manual_seed(0)
    emb = models.ShapeTextureEmbedding(7, 16, 16)
    ids = torch.tensor([3, 3, 3, 5, 5, 0])
    z_s, z_t = emb(ids)
    assert torch.equal(z_s, emb.shape_embedding.weight[ids]) and torch.equal(z_t, emb.texture_embedding.weight[ids])
    src = z_s._cnrf_code_source
    assert src is z_t._cnrf_code_source and src[0] is ids
    assert src[1] is emb.shape_embedding.weight and src[2] is emb.texture_embedding.weight
    assert not hasattr(z_s[:4], "_cnrf_code_source")
    (z_s.sum() + z_t.sum()).backward()                       # the ordinary autograd path still works
    assert emb.shape_embedding.weight.grad[3].eq(3).all() and emb.shape_embedding.weight.grad[1].eq(0).all()
    # the tag records tensor versions: an in-place edit of the codes (or of a table) after the lookup invalidates it
    assert src[3] == (z_s._version, z_t._version, src[1]._version, src[2]._version)
    with torch.no_grad():
        z_s.add_(1.0)
    assert src[3] != (z_s._version, z_t._version, src[1]._version, src[2]._version)


def _ddp_worker(rank, world, port, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import view_synthesis.models as models
    from torch.nn.parallel import DistributedDataParallel as DDP
    torch.manual_seed(0)
    emb = DDP(models.ShapeTextureEmbedding(5, 8, 8))                  # utils/util.py:140, CPU + gloo
    ids = torch.tensor([1, 1, 4])
    z_s, z_t = emb(ids)
    ok = (hasattr(z_s, "_cnrf_code_source") and z_s._cnrf_code_source is z_t._cnrf_code_source
          and z_s._cnrf_code_source[1] is emb.module.shape_embedding.weight)
    if rank == 0:
        torch.save(torch.tensor(int(ok)), out)
    dist.destroy_process_group()


def test_code_source_tag_survives_distributed_data_parallel(tmp_path):
    """train.py --distributed wraps the embedding module in DistributedDataParallel (utils/util.py:139-142): the tag that
    selects the shared-code path must survive DDP.forward's output handling, or every distributed run would silently take
    the per-ray path."""
    out = str(tmp_path / "ok.pt")
    mp.spawn(_ddp_worker, args=(2, 29500 + (os.getpid() % 2000) + 17, out), nprocs=2, join=True)
    assert int(torch.load(out)) == 1


def test_fused_adamw_constructor_and_cpu_refusal():
    """FusedAdamW takes torch.optim.AdamW's arguments, keeps its param-group defaults, refuses the options it does not
    implement and refuses CPU parameters loudly (no fallback)."""
    from codenerf_b200.optim import FusedAdamW
    from codenerf_b200 import _lib
    p = torch.zeros(4, requires_grad=True)
    opt = FusedAdamW([{"params": [p]}], lr=1e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2)
    ref = torch.optim.AdamW([torch.zeros(4, requires_grad=True)], lr=1e-4)
    for k in ("lr", "betas", "eps", "weight_decay"):
        assert opt.param_groups[0][k] == ref.param_groups[0][k], k
    with pytest.raises(ValueError):
        FusedAdamW([p], amsgrad=True)
    with pytest.raises(ValueError):
        FusedAdamW([p], lr=-1.0)
    p.grad = torch.ones(4)
    with pytest.raises(_lib.CnrfError):
        opt.step()
    opt.zero_grad()
    opt.step()                                               # nothing has a gradient: a no-op that needs no GPU


def _grad_worker(rank, world, port, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from codenerf_b200.parallel import GradAllReduce
    torch.manual_seed(100 + rank)                             # ranks start from different parameters, as train.py:29-31 seeds
    lin, emb = torch.nn.Linear(5, 3), torch.nn.Embedding(4, 2)
    frozen = torch.nn.Parameter(torch.ones(2), requires_grad=False)
    sync = GradAllReduce(list(lin.parameters()) + list(emb.parameters()) + [frozen])
    start = [p.detach().clone() for p in lin.parameters()]
    x = torch.full((2, 5), float(rank + 1))
    (lin(x).sum() * (rank + 1)).backward()                    # emb gets no gradient on any rank -> zeros
    sync()
    torch.save({"start": start, "gw": lin.weight.grad.clone(), "gb": lin.bias.grad.clone(), "ge": emb.weight.grad.clone(),
                "flat_views": lin.weight.grad.data_ptr() == sync.flat.data_ptr()}, f"{out}.{rank}")
    dist.destroy_process_group()


def test_grad_allreduce_two_ranks_gloo(tmp_path):
    """GradAllReduce: rank 0's parameters broadcast at construction (as DDP), gradients averaged over ranks through one
    flat buffer, parameters left with views of it, missing gradients treated as zeros, frozen parameters skipped."""
    out = str(tmp_path / "g.pt")
    port = 31500 + (os.getpid() % 2000)
    mp.spawn(_grad_worker, args=(2, port, out), nprocs=2, join=True)
    r0, r1 = torch.load(out + ".0"), torch.load(out + ".1")
    for a, b in zip(r0["start"], r1["start"]):
        assert torch.equal(a, b)                              # broadcast
    # d/dW of (rank+1) * sum(W x + b) with x = rank+1: rank 0 -> 2 rows of 1 * 1, rank 1 -> 2 rows of 2 * 2
    want_w = torch.full((3, 5), (1 * 1 * 2 + 2 * 2 * 2) / 2.0)
    want_b = torch.full((3,), (1 * 2 + 2 * 2) / 2.0)
    for r in (r0, r1):
        assert torch.equal(r["gw"], want_w) and torch.equal(r["gb"], want_b)
        assert torch.equal(r["ge"], torch.zeros(4, 2)) and r["flat_views"]


def test_flat_gradient_allocation_is_16_byte_aligned_per_parameter():
    """ops._flat_grads: one allocation for all parameter gradients of a fused backward, every gradient on a 16-byte boundary
    (fc_out's 257-element bias would otherwise shift everything behind it: the weight-gradient kernel's vector reductions
    and FusedAdamW's float4 path need aligned rows)."""
    import torch
    from codenerf_b200 import ops
    import view_synthesis.models as models
    m = models.CodeNeRFModel(256, 3, 256, 256, 10, 4)
    params = list(m.parameters()) + list(m.parameters())
    flat, grads = ops._flat_grads(params, torch.device("cpu"))
    assert len(grads) == len(params) and flat.dim() == 1
    end = 0
    for g, p in zip(grads, params):
        assert g.shape == p.shape and g.is_contiguous()
        assert g.storage_offset() % 4 == 0 and g.storage_offset() >= end          # aligned, in order, no overlap
        assert g.untyped_storage().data_ptr() == flat.untyped_storage().data_ptr()
        end = g.storage_offset() + g.numel()
    assert end <= flat.numel() < end + 4


def test_loss_target_spec_argument_checks():
    """ops.LossTargetSpec: exactly one of (rows) / (uint8 views + int64 pixel indices); CPU tensors are refused when the
    struct is built (there is no CPU fallback)."""
    import pytest
    import torch
    from codenerf_b200 import ops
    with pytest.raises(AssertionError):
        ops.LossTargetSpec()
    with pytest.raises(AssertionError):
        ops.LossTargetSpec(rows=torch.zeros(4, 3), views=torch.zeros(1, 2, 2, 4, dtype=torch.uint8), pixel=torch.zeros(4, dtype=torch.int64))
    with pytest.raises(AssertionError):
        ops.LossTargetSpec(views=torch.zeros(1, 2, 2, 4), pixel=torch.zeros(4, dtype=torch.int64))       # float views
    spec = ops.LossTargetSpec(views=torch.zeros(1, 2, 2, 4, dtype=torch.uint8), pixel=torch.zeros(4, dtype=torch.int64))
    with pytest.raises(AssertionError):
        spec.struct(4, [])                                                                             # CPU tensors
