"""No-grad CodeNeRF rendering throughput (parallel_image_render's inner loop, nerf/__init__.py:195-205): 128x128 views of one
object (stride-0 latent codes, eval.py:149), 32 + 128 samples, validation chunks of 4096 rays, bf16 MLP; rays/s by CUDA events.

    python scripts/time_codenerf_render.py [views=16] [out.json]
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

views = int(sys.argv[1]) if len(sys.argv) > 1 else 16
out_path = sys.argv[2] if len(sys.argv) > 2 else None
dev = torch.device("cuda", 0)
torch.manual_seed(0)
nerf.set_precision("bf16")
coarse = models.CodeNeRFModel(256, 1, 256, 256, 10, 4).to(dev)
fine = models.CodeNeRFModel(256, 1, 256, 256, 10, 4).to(dev)
ps = nerf.PointSampler(32, 128, 0.8, 1.8, "lindepth", True, torch.float32, dev)
emb = (nerf.PositionalEmbedder(10, True, True, torch.float32, dev), nerf.PositionalEmbedder(4, True, True, torch.float32, dev))
pose = torch.eye(4, device=dev)[None].clone()
pose[0, 2, 3] = 1.3
_, rd = ops.ray_bundle(pose, 128, 128, 131.25, 64.0, 64.0, want_ro=False)
ro = pose[0, :3, 3].expand(rd.shape[0], 3).contiguous()
zs, zt = torch.randn(1, 256, device=dev), torch.randn(1, 256, device=dev)
chunk = 4096


def render():
    outs = []
    with torch.no_grad():
        for i in range(0, rd.shape[0], chunk):
            n = min(chunk, rd.shape[0] - i)
            _, rgb = nerf.predict_radiance_and_render((ro[i:i + n], rd[i:i + n]), ps, emb, coarse, fine,
                                                      (zs.expand(n, -1), zt.expand(n, -1)))
            outs.append(rgb)
    return torch.cat(outs)


for _ in range(2):
    render()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(views):
    img = render()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
rays = views * rd.shape[0]
assert torch.isfinite(img).all()
res = {"workload": "CodeNeRF 128x128 no-grad render, 32+128 samples, chunks of 4096, bf16", "views": views, "rays": rays, "ms": ms,
       "rays_per_s": rays / (ms * 1e-3), "ms_per_view": ms / views}
print(json.dumps(res))
if out_path:
    json.dump(res, open(out_path, "w"), indent=1)
