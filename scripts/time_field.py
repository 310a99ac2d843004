"""Time the field forward (and backward when available) at C1 shapes; prints ms and TFLOP/s."""
import os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "code-nerf_b200")); sys.path.insert(0, ROOT)
import view_synthesis.nerf as nerf, view_synthesis.models as models
from codenerf_b200 import _lib
dev = torch.device("cuda", 0)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
S = int(sys.argv[2]) if len(sys.argv) > 2 else 160
mode = sys.argv[3] if len(sys.argv) > 3 else "fwd"
prec = sys.argv[4] if len(sys.argv) > 4 else "bf16"
nerf.set_precision(prec)
m = models.CodeNeRFModel(256, 1, 256, 256, 10, 4, True, True).to(dev)
emb = (nerf.PositionalEmbedder(10, True, True, torch.float32, dev), nerf.PositionalEmbedder(4, True, True, torch.float32, dev))
pts, rd = torch.randn(N, S, 3, device=dev), torch.randn(N, 3, device=dev)
zs, zt = torch.randn(N, 256, device=dev, requires_grad=(mode != "fwd")), torch.randn(N, 256, device=dev)
FS = 2 * (63 * 256 + 256 * 256 + 256 * 257 + 256 * 256 + 256 * 256 + 256 * 3)
def step():
    if mode == "fwd":
        with torch.no_grad():
            return nerf.forward_pass(m, emb, rd, pts, (zs, zt))
    raw = nerf.forward_pass(m, emb, rd, pts, (zs, zt))
    raw.backward(torch.ones_like(raw))
for _ in range(3): step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 10
e0.record()
for _ in range(reps): step()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
fl = N * S * FS * (1 if mode == "fwd" else 3)
print(f"{prec} {mode} N={N} S={S}: {ms:.3f} ms  {fl / ms / 1e9:.1f} TFLOP/s  tc_err={_lib.load().cnrf_debug_tc_error()}")
try:
    import numpy as np, ctypes
    lib = _lib.load()
    buf = np.zeros(16, dtype=np.int64)
    lib.cnrf_debug_fwd_times(ctypes.c_void_p(buf.ctypes.data), 1)
    step(); torch.cuda.synchronize()
    lib.cnrf_debug_fwd_times(ctypes.c_void_p(buf.ctypes.data), 1)
    n = max(1, int(buf[7]))
    print("block0/tile0 per pair [clk]: wait_acc(total of 5) %d | epi L0..L4 %s | build %d | MMA-warp waits: weights %d, a_ready %d | pairs %d" %
          (buf[0] // n, [int(buf[1 + i] // n) for i in range(5)], buf[6] // n, buf[8] // n, buf[9] // n, n))
    print("  epilogue split per pair (5 layers): store-drain + barrier %d | blocks %d | closing barrier %d | store issue + arrive %d" %
          (buf[10] // n, buf[11] // n, buf[12] // n, buf[13] // n))
    print("  MMA warp per pair: fence + 4 MMA issues (sum over images) %d | commits %d" % (buf[14] // n, buf[15] // n))
    tr_all = np.zeros(128, dtype=np.int64)
    lib.cnrf_debug_fwd_trace(ctypes.c_void_p(tr_all.ctypes.data))
    tr, ti = tr_all[:64].reshape(2, 4, 8), tr_all[64:].reshape(16, 4)
    if tr.any():
        t0 = tr[tr > 0].min()
        names = ["wait_acc", "acc_seen", "blocks_start", "blocks_done", "handoff", "mma_sees_a_ready", "mma_issued", "gamma_built"]
        ev = sorted((int(tr[t, l, e] - t0), f"T{t} L{l} {names[e]}") for t in range(2) for l in range(4) for e in range(8) if tr[t, l, e] > 0)
        print("timeline of block 0's 4th tile pair [clk from first event]:")
        for clk, what in ev:
            print(f"  {clk:7d}  {what}")
        if ti.any():
            print("weight images of that pair's layer 1 (T0: 0-7, T1: 8-15) [clk from the pair's first event]:")
            print("  img  slot_empty_seen  copy_issued  image_seen_by_MMA  MMAs_issued   (issue->seen)")
            for i in range(16):
                a, b, c, d = (int(x - t0) if x > 0 else -1 for x in ti[i])
                print(f"  {i:3d}  {a:15d}  {b:11d}  {c:17d}  {d:11d}   {c - b:8d}")
except Exception as e:
    print("no fwd debug times:", e)
