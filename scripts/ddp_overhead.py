"""Where the data-parallel step's extra time goes (2+ ranks): the bench's chunk step with DDP as utils/util.py:139-142 wraps
the modules, (a) as is, (b) under no_sync() (no gradient all-reduce: host-side DDP bookkeeping only), (c) host enqueue time.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/ddp_overhead.py
"""
import contextlib
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "code-nerf_b200")); sys.path.insert(0, ROOT)
import view_synthesis.nerf as nerf, view_synthesis.models as models  # noqa: E402
from codenerf_b200.optim import FusedAdamW  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
torch.manual_seed(rank); np.random.seed(rank)
N = 4096
raw = {"embedding": models.ShapeTextureEmbedding(2458, 256, 256).to(dev),
       "nerf_coarse": models.CodeNeRFModel(256, 2458, 256, 256, 10, 4).to(dev),
       "nerf_fine": models.CodeNeRFModel(256, 2458, 256, 256, 10, 4).to(dev)}
from torch.nn.parallel import DistributedDataParallel as DDP
mods = {k: DDP(m, device_ids=[local], output_device=local) for k, m in raw.items()}
opt = FusedAdamW([{"params": mods["nerf_coarse"].parameters()}, {"params": mods["nerf_fine"].parameters()},
                  {"params": mods["embedding"].parameters(), "lr": 1e-3}], lr=1e-4)
ps = nerf.PointSampler(32, 128, 0.8, 1.8, "lindepth", True, torch.float32, dev)
emb = (nerf.PositionalEmbedder(10, True, True, torch.float32, dev), nerf.PositionalEmbedder(4, True, True, torch.float32, dev))
ro = torch.tensor([0.0, 0.0, 1.3], device=dev).expand(N, 3).contiguous() + 0.01 * torch.randn(N, 3, device=dev)
rd = torch.nn.functional.normalize(torch.randn(N, 3, device=dev) * 0.3 + torch.tensor([0.0, 0.0, -1.0], device=dev), dim=-1)
tgt = torch.rand(N, 3, device=dev)
oid = torch.full((N,), 1234, dtype=torch.int64, device=dev)


def step(use):
    codes = use["embedding"](oid)
    a, b = nerf.predict_radiance_and_render((ro, rd), ps, emb, use["nerf_coarse"], use["nerf_fine"], codes)
    loss = torch.nn.functional.mse_loss(a, tgt) + torch.nn.functional.mse_loss(b, tgt)
    opt.zero_grad()
    loss.backward()
    opt.step()


def timed(fn, reps=40):
    for _ in range(5):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    host = (time.perf_counter() - t0) / reps * 1e3
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, host


def nosync_step():
    with contextlib.ExitStack() as st:
        for m in mods.values():
            st.enter_context(m.no_sync())
        step(mods)


res = {"ddp": timed(lambda: step(mods)), "ddp_no_sync": timed(nosync_step), "plain_modules": timed(lambda: step(raw))}
if rank == 0:
    for k, (gpu, host) in res.items():
        print(f"{k:14s}: chunk {gpu:.3f} ms on the GPU timeline, host enqueue {host:.3f} ms")
dist.destroy_process_group()
