"""Forward-chain kernel time with and without the saved images (inference vs training forward), fine-pass shape."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "code-nerf_b200")); sys.path.insert(0, ROOT)
import view_synthesis.nerf as nerf, view_synthesis.models as models
from codenerf_b200 import _lib
dev = torch.device("cuda", 0)
N, S = 4096, 160
nerf.set_precision("bf16")
m = models.CodeNeRFModel(256, 1, 256, 256, 10, 4, True, True).to(dev)
emb = (nerf.PositionalEmbedder(10, True, True, torch.float32, dev), nerf.PositionalEmbedder(4, True, True, torch.float32, dev))
pts, rd = torch.randn(N, S, 3, device=dev), torch.randn(N, 3, device=dev)
zs, zt = torch.randn(N, 256, device=dev), torch.randn(N, 256, device=dev)
lib = _lib.load()
for mode in ("inference", "training"):
    def step():
        if mode == "inference":
            with torch.no_grad():
                return nerf.forward_pass(m, emb, rd, pts, (zs, zt))
        return nerf.forward_pass(m, emb, rd, pts, (zs.requires_grad_(True), zt))
    for _ in range(3): step()
    torch.cuda.synchronize()
    lib.cnrf_debug_kernel_ms(0, N * S, None, None)
    for _ in range(10): step()
    torch.cuda.synchronize()
    print(f"{mode}: forward-chain kernel {lib.cnrf_debug_kernel_ms(0, 0, None, None) * 1e3:.0f} us  (pairs={os.environ.get('CNRF_FWD_CTA_PAIRS', '1')})")
