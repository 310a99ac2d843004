"""GPU-busy time vs wall time of the training chunk (torch.profiler, CUPTI kernel records)."""
import os, sys, torch, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "code-nerf_b200")); sys.path.insert(0, ROOT)
import view_synthesis.nerf as nerf, view_synthesis.models as models
from torch.profiler import profile, ProfilerActivity
N = 4096
dev = torch.device("cuda", 0)
nerf.set_precision("bf16")
torch.manual_seed(0)
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
oid = torch.full((N,), 1234, dtype=torch.int64, device=dev)
def step():
    codes = mods["embedding"](oid)
    a, b = nerf.predict_radiance_and_render((ro, rd), ps, emb, mods["nerf_coarse"], mods["nerf_fine"], codes)
    loss = torch.nn.functional.mse_loss(a, tgt) + torch.nn.functional.mse_loss(b, tgt)
    opt.zero_grad(); loss.backward(); opt.step()
for _ in range(5): step()
torch.cuda.synchronize()
reps = 8
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): step()
    e1.record(); torch.cuda.synchronize()
wall = e0.elapsed_time(e1) / reps
rows = [(e.key, e.device_time_total / reps, e.count // reps) for e in prof.key_averages() if e.device_time_total > 0]
rows.sort(key=lambda r: -r[1])
busy = sum(r[1] for r in rows) / 1e3
print(f"chunk wall {wall:.3f} ms, GPU busy {busy:.3f} ms ({100 * busy / wall:.1f}%), {sum(r[2] for r in rows)} kernels+memops")
for k, t, n in rows[:14]:
    print(f"  {t:8.1f} us x{n:3d}  {k[:100]}")
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "gemm_tc" in e.name]
print("gemm_tc launches of the last chunk [us]:", [round(e.device_time, 1) for e in ev[-8:]])
kern = {}
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA:
        k = kern.setdefault(e.name, [0.0, 0]); k[0] += e.device_time / reps; k[1] += 1
tot = sum(v[0] for v in kern.values())
print(f"device kernels/memops only: {tot / 1e3:.3f} ms per chunk")
for name, (t, n) in sorted(kern.items(), key=lambda kv: -kv[1][0])[:40]:
    print(f"  {t:8.1f} us x{n // reps:3d}  {name[:110]}")
