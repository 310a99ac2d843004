"""Host (enqueue) time of the pieces of one training chunk: how long Python + the launch path take, GPU running async."""
import os, sys, time, torch, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "code-nerf_b200")); sys.path.insert(0, ROOT)
import view_synthesis.nerf as nerf, view_synthesis.models as models
from codenerf_b200.optim import FusedAdamW
N = 4096
dev = torch.device("cuda", 0)
nerf.set_precision("bf16")
mods = {"embedding": models.ShapeTextureEmbedding(2458, 256, 256).to(dev),
        "nerf_coarse": models.CodeNeRFModel(256, 2458, 256, 256, 10, 4).to(dev),
        "nerf_fine": models.CodeNeRFModel(256, 2458, 256, 256, 10, 4).to(dev)}
opt = FusedAdamW([{"params": mods["nerf_coarse"].parameters()}, {"params": mods["nerf_fine"].parameters()},
                  {"params": mods["embedding"].parameters(), "lr": 1e-3}], lr=1e-4)
ps = nerf.PointSampler(32, 128, 0.8, 1.8, "lindepth", True, torch.float32, dev)
emb = (nerf.PositionalEmbedder(10, True, True, torch.float32, dev), nerf.PositionalEmbedder(4, True, True, torch.float32, dev))
ro = torch.tensor([0.0, 0.0, 1.3], device=dev).expand(N, 3).contiguous()
rd = torch.nn.functional.normalize(torch.randn(N, 3, device=dev) * 0.3 + torch.tensor([0.0, 0.0, -1.0], device=dev), dim=-1)
tgt = torch.rand(N, 3, device=dev)
oid = torch.full((N,), 1234, dtype=torch.int64, device=dev)
T = {}
def tick(name, t0):
    T[name] = T.get(name, 0.0) + time.perf_counter() - t0
def step():
    t = time.perf_counter(); codes = mods["embedding"](oid); tick("embedding", t)
    t = time.perf_counter()
    a, b = nerf.predict_radiance_and_render((ro, rd), ps, emb, mods["nerf_coarse"], mods["nerf_fine"], codes); tick("render fwd", t)
    t = time.perf_counter(); loss = torch.nn.functional.mse_loss(a, tgt) + torch.nn.functional.mse_loss(b, tgt); tick("loss", t)
    t = time.perf_counter(); opt.zero_grad(); tick("zero_grad", t)
    t = time.perf_counter(); loss.backward(); tick("backward", t)
    t = time.perf_counter(); opt.step(); tick("opt.step", t)
for _ in range(5): step()
torch.cuda.synchronize(); T.clear()
reps = 30
for _ in range(reps):
    step()
    torch.cuda.synchronize()        # so that the host never blocks on a full launch queue
print("host ms per chunk:", {k: round(v / reps * 1e3, 3) for k, v in T.items()}, "total", round(sum(T.values()) / reps * 1e3, 3))
