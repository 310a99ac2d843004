"""Mean CUDA-event duration of the three tensor-core kernels (forward chain, backward-data chain, weight gradients) over
the fine-pass launches of repeated 4096-ray training chunks.

    [CNRF_DW_JOBMASK=m] python scripts/time_kernels.py [rays=4096] [reps=10]
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
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
dev = torch.device("cuda", 0)
nerf.set_precision("bf16")
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
lib = _lib.load()


def step():
    codes = mods["embedding"](oid)
    a, b = nerf.predict_radiance_and_render((ro, rd), ps, emb, mods["nerf_coarse"], mods["nerf_fine"], codes)
    _, _, loss = ops.mse_pair(a, b, tgt)          # bench.py's loss (train.py:103-104 in one launch)
    opt.zero_grad()
    loss.backward()
    opt.step()


for _ in range(3):
    step()
torch.cuda.synchronize()
fine_rows = N * 160
for k in range(3):
    lib.cnrf_debug_kernel_ms(k, fine_rows, None, None)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
import time
t_host0 = time.perf_counter()
e0.record()
for _ in range(reps):
    step()
e1.record()
t_host = (time.perf_counter() - t_host0) / reps * 1e3       # host time to ENQUEUE one chunk (no synchronisation)
torch.cuda.synchronize()
ms = [lib.cnrf_debug_kernel_ms(k, 0, None, None) for k in range(3)]
print(f"mask={os.environ.get('CNRF_DW_JOBMASK', 'all')} chunk {e0.elapsed_time(e1) / reps:.3f} ms (host enqueue {t_host:.3f} ms) | fine launches: fwd {ms[0] * 1e3:.0f} us, "
      f"bwd {ms[1] * 1e3:.0f} us, dW {ms[2] * 1e3:.0f} us | tc_err={lib.cnrf_debug_tc_error()}")
