"""In-stream timing of the per-ray GEMM toolkit (tc=1 tcgen05 split-bf16, tc=0 fp32 SIMT)."""
import ctypes as C, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "code-nerf_b200"))
from codenerf_b200 import _lib
lib = _lib.load(); dev = torch.device("cuda", 0)
def run(tc, M, N, K, form, reps=50):
    X, W, dY = torch.randn(M, K, device=dev), torch.randn(N, K, device=dev), torch.randn(M, N, device=dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream); p = lambda t: C.c_void_p(t.data_ptr())
    if form == "lin":   out = torch.zeros(M, N, device=dev); f = lambda: lib.cnrf_debug_gemm(tc, p(X), K, 1, p(W), 1, K, p(out), N, M, N, K, 0, st)
    if form == "dgrad": out = torch.zeros(M, K, device=dev); f = lambda: lib.cnrf_debug_gemm(tc, p(dY), N, 1, p(W), K, 1, p(out), K, M, K, N, 0, st)
    if form == "wgrad": out = torch.zeros(N, K, device=dev); f = lambda: lib.cnrf_debug_gemm(tc, p(dY), 1, N, p(X), K, 1, p(out), K, N, K, M, 1, st)
    for _ in range(5): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    print(f"tc={tc} {form:6s} M={M} N={N} K={K}: {e0.elapsed_time(e1) / reps * 1e3:.1f} us")
for tc in (1, 0):
    for form in ("lin", "dgrad", "wgrad"):
        run(tc, 4096, 256, 256, form)
    run(tc, 4096, 256, 27, "lin"); run(tc, 4096, 3, 256, "lin"); run(tc, 128, 64, 16, "lin")

