"""Timing of ONE big per-sample GEMM of the toolkit (the shapes of vanilla-NeRF training / the split pose chain):
    python scripts/time_gemm_big.py [M=786432]          CNRF_GEMM_NO_STREAM=1 selects the batched kernel's wide tiles"""
import ctypes as C, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "code-nerf_b200"))
from codenerf_b200 import _lib
lib = _lib.load(); dev = torch.device("cuda", 0)
M = int(sys.argv[1]) if len(sys.argv) > 1 else 786432
def run(N, K, form, reps=10):
    X, W, dY = torch.randn(M, K, device=dev), torch.randn(N, K, device=dev), torch.randn(M, N, device=dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream); p = lambda t: C.c_void_p(t.data_ptr())
    if form == "lin":   out = torch.zeros(M, N, device=dev); f = lambda: lib.cnrf_debug_gemm(1, p(X), K, 1, p(W), 1, K, p(out), N, M, N, K, 0, st)
    if form == "dgrad": out = torch.zeros(M, K, device=dev); f = lambda: lib.cnrf_debug_gemm(1, p(dY), N, 1, p(W), K, 1, p(out), K, M, K, N, 0, st)
    if form == "wgrad": out = torch.zeros(N, K, device=dev); f = lambda: lib.cnrf_debug_gemm(1, p(dY), 1, N, p(X), K, 1, p(out), K, N, K, M, 1, st)
    for _ in range(2): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / reps * 1e3
    if form == "wgrad":
        out.zero_(); f(); torch.cuda.synchronize()
        ref = dY.double().t() @ X.double()
        err = float((out.double() - ref).abs().max() / ref.abs().max())
        print(f"{form:6s} M={M} N={N} K={K}: {us:.1f} us  {2 * M * N * K / us / 1e6:.0f} TFLOP/s  {(M * K + M * N) * 4 / 1e9 / us * 1e6:.0f} GB/s  max rel err {err:.1e}")
        return
    ref = (X @ W.t()) if form == "lin" else (dY @ W)
    err = float((out - ref).abs().max() / ref.abs().max())
    gb = (M * K + M * N) * 4 / 1e9 if form == "lin" else (M * N + M * K) * 4 / 1e9
    print(f"{form:6s} M={M} N={N} K={K}: {us:.1f} us  {2 * M * N * K / us / 1e6:.0f} TFLOP/s  {gb / us * 1e6:.0f} GB/s  max rel err {err:.1e}")
run(256, 64, "lin"); run(128, 256, "lin"); run(256, 256, "dgrad"); run(256, 256, "wgrad"); run(256, 256, "lin")
buf = (C.c_longlong * 64)()
lib.cnrf_debug_stream_trace(C.cast(buf, C.c_void_p))
tr = list(buf)
if any(tr):
    t0 = min(v for v in tr if v > 0)
    r = lambda i: (tr[i] - t0) if tr[i] > 0 else -1
    print("stream trace (last call, block 0, 6th tile) [clk]:")
    for ch in range(5):
        if tr[3 * ch] > 0:
            print(f"  chunk {ch}: conv loads_issued {r(3*ch)} slot_free {r(3*ch+1)} arrived {r(3*ch+2)} | mma A {r(16+3*ch)} B {r(17+3*ch)} issued {r(18+3*ch)} | tma free {r(36+2*ch)} issued {r(37+2*ch)}")
    print(f"  mma acc_free {r(31)} committed {r(32)} | epi w0 seen {r(48)} done {r(49)} | epi w4 seen {r(50)} done {r(51)}")
    if tr[52] > 0:
        print(f"  epi w0 block 2: start {r(52)} tmem_wait {tr[53]-tr[52]} stage+sync {tr[54]-tr[53]} turn+store {tr[55]-tr[54]} sync {tr[56]-tr[55]}")
