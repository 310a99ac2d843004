"""BASELINE config 5 (config/lego.yml) TRAINING step shape: 16384 random rays per step in chunks of 4096 (nerf.train.chunksize),
FlexibleNeRFModel 8 x 256 skip [4] coarse + fine, 64 + 128 samples, forward + backward + Adam through the drop-in
predict_radiance_and_render; rays/s by CUDA events.  `keep = 1` runs the per-layer toolkit GEMMs (csrc/flexnerf.cu).

    python scripts/time_lego_train.py [steps=3] [out.json]
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "code-nerf_b200")); sys.path.insert(0, ROOT)
import view_synthesis.nerf as nerf  # noqa: E402
import view_synthesis.models as models  # noqa: E402
from codenerf_b200.optim import FusedAdamW  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
out_path = sys.argv[2] if len(sys.argv) > 2 else None
dev = torch.device("cuda", 0)
torch.manual_seed(0)
N, chunk = 16384, 4096
coarse = models.FlexibleNeRFModel(8, 256, [4], 10, 4).to(dev)
fine = models.FlexibleNeRFModel(8, 256, [4], 10, 4).to(dev)
opt = FusedAdamW(list(coarse.parameters()) + list(fine.parameters()), lr=5e-3, weight_decay=0.0)
ps = nerf.PointSampler(64, 128, 2.0, 6.0, "lindepth", True, torch.float32, dev)
emb = (nerf.PositionalEmbedder(10, True, True, torch.float32, dev), nerf.PositionalEmbedder(4, True, True, torch.float32, dev))
ro = torch.tensor([0.0, 0.0, 4.0], device=dev).expand(N, 3).contiguous()
rd = torch.nn.functional.normalize(torch.randn(N, 3, device=dev) * 0.3 + torch.tensor([0.0, 0.0, -1.0], device=dev), dim=-1)
tgt = torch.rand(N, 3, device=dev)
res = {"workload": "lego.yml train step: 16384 rays in chunks of 4096, FlexibleNeRFModel 8x256 skip[4] x2, 64+128 samples, fwd+bwd+Adam"}
for precision in ("bf16", "fp32"):
    nerf.set_precision(precision)

    def step():
        opt.zero_grad(set_to_none=True)
        for i in range(0, N, chunk):
            rc, rf = nerf.predict_radiance_and_render((ro[i:i + chunk], rd[i:i + chunk]), ps, emb, coarse, fine, (None, None))
            loss = ((rc - tgt[i:i + chunk]) ** 2).mean() + ((rf - tgt[i:i + chunk]) ** 2).mean()
            (loss * (chunk / N)).backward()
        opt.step()
        return loss
    n = steps if precision == "bf16" else 1
    step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    assert torch.isfinite(loss)
    # 1,186,816 FLOP per sample forward, x3 for forward + both backward products, 256 sample evaluations per ray
    res[precision] = {"ms_per_step": ms, "rays_per_s": N / (ms * 1e-3), "tflops_algorithmic": N * 256 * 1186816 * 3 / (ms * 1e-3) / 1e12}
    print(f"{precision}: {ms:.1f} ms per step = {N / (ms * 1e-3):,.0f} rays/s ({res[precision]['tflops_algorithmic']:.1f} algorithmic TFLOP/s)")
if out_path:
    json.dump(res, open(out_path, "w"), indent=1)
