"""One training chunk (4096 rays, coarse+fine fwd + 2 MSE + backward + AdamW) of the C2 workload, repeated; the last
repetition sits between cudaProfilerStart/Stop so `ncu --profile-from-start off` sees exactly one chunk step.

    python scripts/profile_chunk.py [rays=4096] [reps=3] [precision=bf16]
"""
import os
import sys
import torch
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "code-nerf_b200")); sys.path.insert(0, ROOT)
import view_synthesis.nerf as nerf, view_synthesis.models as models
from codenerf_b200 import _lib, ops

N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
prec = sys.argv[3] if len(sys.argv) > 3 else "bf16"
dev = torch.device("cuda", 0)
nerf.set_precision(prec)
torch.manual_seed(0); np.random.seed(0)
mods = {"embedding": models.ShapeTextureEmbedding(2458, 256, 256).to(dev),
        "nerf_coarse": models.CodeNeRFModel(256, 2458, 256, 256, 10, 4).to(dev),
        "nerf_fine": models.CodeNeRFModel(256, 2458, 256, 256, 10, 4).to(dev)}
from codenerf_b200.optim import FusedAdamW
opt = FusedAdamW([{"params": mods["nerf_coarse"].parameters()}, {"params": mods["nerf_fine"].parameters()},
                         {"params": mods["embedding"].parameters(), "lr": 1e-3}], lr=1e-4)
ps = nerf.PointSampler(32, 128, 0.8, 1.8, "lindepth", True, torch.float32, dev)
emb = (nerf.PositionalEmbedder(10, True, True, torch.float32, dev), nerf.PositionalEmbedder(4, True, True, torch.float32, dev))
ro = torch.tensor([0.0, 0.0, 1.3], device=dev).expand(N, 3).contiguous() + 0.01 * torch.randn(N, 3, device=dev)
rd = torch.nn.functional.normalize(torch.randn(N, 3, device=dev) * 0.3 + torch.tensor([0.0, 0.0, -1.0], device=dev), dim=-1)
tgt = torch.rand(N, 3, device=dev)
oid = torch.full((N,), 1234, dtype=torch.int64, device=dev) if os.environ.get("CHUNK_IDS", "image") == "image" else torch.randint(0, 2458, (N,), device=dev)   # train.py:78: one object id per image


def step():
    codes = mods["embedding"](oid)
    a, b = nerf.predict_radiance_and_render((ro, rd), ps, emb, mods["nerf_coarse"], mods["nerf_fine"], codes)
    _, _, loss = ops.mse_pair(a, b, tgt)          # bench.py's loss (train.py:103-104 in one launch)
    opt.zero_grad()
    loss.backward()
    opt.step()
    return loss


for _ in range(reps - 1):
    step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.profiler.start()
e0.record()
loss = step()
e1.record()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print(f"{prec} chunk step N={N}: {e0.elapsed_time(e1):.3f} ms  loss={loss.item():.5f}  tc_err={_lib.load().cnrf_debug_tc_error()}")
