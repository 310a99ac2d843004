"""BASELINE config 5 (config/lego.yml): forward-only render of one 800x800 view (640,000 rays) with the vanilla
NeRF network (FlexibleNeRFModel 8 x 256, skip [4]), 64 coarse + 128 fine samples per ray, near/far 2/6, validation chunks of
8192 rays, through the drop-in predict_radiance_and_render; rays/s by CUDA events, both MLP precisions.

    python scripts/time_lego_render.py [rays=640000] [out.json]
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "code-nerf_b200")); sys.path.insert(0, ROOT)
import view_synthesis.nerf as nerf  # noqa: E402
import view_synthesis.models as models  # noqa: E402
from codenerf_b200 import ops  # noqa: E402

n_rays = int(sys.argv[1]) if len(sys.argv) > 1 else 640000
out_path = sys.argv[2] if len(sys.argv) > 2 else None
dev = torch.device("cuda", 0)
torch.manual_seed(0)
Hh = Ww = 800
chunk = 8192
coarse = models.FlexibleNeRFModel(8, 256, [4], 10, 4).to(dev)
fine = models.FlexibleNeRFModel(8, 256, [4], 10, 4).to(dev)
ps = nerf.PointSampler(64, 128, 2.0, 6.0, "lindepth", True, torch.float32, dev)
emb = (nerf.PositionalEmbedder(10, True, True, torch.float32, dev), nerf.PositionalEmbedder(4, True, True, torch.float32, dev))
pose = torch.eye(4, device=dev)[None].clone()
pose[0, 2, 3] = 4.0
_, rd = ops.ray_bundle(pose, Hh, Ww, 1111.0, Ww / 2, Hh / 2, want_ro=False)
rd = rd[:n_rays].contiguous()
ro = pose[0, :3, 3].expand(rd.shape[0], 3).contiguous()
res = {"workload": "lego.yml render: FlexibleNeRFModel 8x256 skip[4], 64+128 samples, chunks of 8192 rays", "rays": int(rd.shape[0])}
for precision in ("bf16", "split", "fp32"):
    nerf.set_precision(precision)
    # fp32 = the CUDA-core parity mode; split = the same per-layer path on split-bf16 tensor-core GEMMs (fp32-grade accuracy)
    rays = rd.shape[0] if precision == "bf16" else min(rd.shape[0], (32 if precision == "split" else 8) * chunk)

    def render():
        outs = []
        with torch.no_grad():
            for i in range(0, rays, chunk):
                _, rgb = nerf.predict_radiance_and_render((ro[i:i + chunk], rd[i:i + chunk]), ps, emb, coarse, fine, (None, None))
                outs.append(rgb)
        return torch.cat(outs)
    render() if precision == "fp32" else [render() for _ in range(1)]
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    img = render()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    assert torch.isfinite(img).all()
    res[precision] = {"rays": rays, "ms": ms, "rays_per_s": rays / (ms * 1e-3),
                      "tflops_algorithmic": rays * 303.8e6 / (ms * 1e-3) / 1e12}
    print(f"{precision}: {rays} rays in {ms:.1f} ms = {rays / (ms * 1e-3):,.0f} rays/s "
          f"({rays * 303.8e6 / (ms * 1e-3) / 1e12:.1f} algorithmic TFLOP/s)")
if out_path:
    json.dump(res, open(out_path, "w"), indent=1)
