"""Timeline of one steady-state tile of the fused vanilla-NeRF chain (flex_fwd_tc.cu), from a -DCNRF_FLEX_TIMERS build:

    CNRF_NVCC_EXTRA=-DCNRF_FLEX_TIMERS python -c "import sys; sys.path.insert(0, 'code-nerf_b200'); import build; build.build(force=True)"
    python scripts/trace_flex.py [out.txt]

Prints, per layer and relative to the tile's first event, when the epilogue saw the accumulator and finished each slab,
when the MMA warp saw each A slab / weight slab and issued its MMAs, and when the producer saw a free slot / issued the copy."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "code-nerf_b200")); sys.path.insert(0, ROOT)
import view_synthesis.nerf as nerf  # noqa: E402
import view_synthesis.models as models  # noqa: E402
from codenerf_b200 import _lib  # noqa: E402

dev = torch.device("cuda", 0)
torch.manual_seed(0)
m = models.FlexibleNeRFModel(8, 256, [4], 10, 4).to(dev)
emb = (nerf.PositionalEmbedder(10, True, True, torch.float32, dev), nerf.PositionalEmbedder(4, True, True, torch.float32, dev))
N, S = 8192, 192
rd, pts = torch.randn(N, 3, device=dev), torch.rand(N, S, 3, device=dev) * 4 - 2
nerf.set_precision("bf16")
with torch.no_grad():
    for _ in range(2):
        nerf.forward_pass(m, emb, rd, pts, (None, None))
torch.cuda.synchronize()
buf = (C.c_longlong * 720)()
_lib.check(_lib.load().cnrf_debug_flex_trace(C.cast(buf, C.c_void_p)), "cnrf_debug_flex_trace")
tr = [[buf[l * 40 + e] for e in range(40)] for l in range(18)]
t0 = min(v for row in tr for v in row if v > 0) if any(v > 0 for row in tr for v in row) else 0
lines = []
if t0 == 0:
    lines.append("no stamps: not a -DCNRF_FLEX_TIMERS build")
rel = lambda v: f"{v - t0:7d}" if v > 0 else "      -"
for l in range(10):
    r = tr[l]
    lines.append(f"layer {l}: epi acc_seen {rel(r[0])} slabs {' '.join(rel(r[1 + i]) for i in range(4))}" + (f" gamma {rel(r[5])}" if l == 0 else ""))
    lines.append("         mma " + " | ".join(f"A{rel(r[8 + 3 * i])} W{rel(r[9 + 3 * i])} iss{rel(r[10 + 3 * i])}" for i in range(5) if r[10 + 3 * i] > 0) + f" commit {rel(r[23])}")
    if r[30] > 0:
        lines.append(f"         epi slab 1 of warp 0: wait_ld {r[31] - r[30]} convert+store {r[32] - r[31]} fence.proxy.async {r[33] - r[32]} arrive {r[2] - r[33]}")
    if l == 1 and r[34] > 0:
        lines.append(f"         gamma(d) build of warp 0: start {rel(r[34])} loads issued {rel(r[35])} stored {rel(r[36])} arrived {rel(r[37])}")
    lines.append("         tma " + " | ".join(f"empty{rel(r[24 + 2 * i])} iss{rel(r[25 + 2 * i])}" for i in range(5) if r[25 + 2 * i] > 0))
txt = "\n".join(lines)
print(txt)
if len(sys.argv) > 1:
    open(sys.argv[1], "w").write(txt + "\n")
