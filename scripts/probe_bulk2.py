"""Where the ~587 clk per cp.async.bulk copy comes from (csrc/tc_probe.cu: cnrf_debug_bulk_ld2).
    python scripts/probe_bulk2.py > gpurun_out/bulk_probe2.json"""
import ctypes as C, json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "code-nerf_b200"))
from codenerf_b200 import _lib  # noqa: E402
lib = _lib.load()
dev = torch.device("cuda", 0)
src = torch.zeros(2 * 1024 * 1024, dtype=torch.uint8, device=dev)
rows = []
for blocks in (1, 148):
    for mode, nthreads, nbytes, slots in ((1, 1, 16384, 4), (1, 2, 16384, 4), (1, 4, 16384, 3), (1, 4, 8192, 4), (2, 1, 16384, 4), (2, 1, 16384, 8),
                                          (2, 1, 4096, 16), (2, 1, 32768, 4), (3, 1, 16384, 4), (3, 1, 16384, 8), (3, 1, 8192, 16)):
        iters = 1920
        clocks = torch.zeros(blocks, dtype=torch.int64, device=dev)
        for _ in range(2):
            _lib.check(lib.cnrf_debug_bulk_ld2(blocks, slots, nbytes, iters, mode, nthreads, C.c_void_p(src.data_ptr()), src.numel(),
                                               C.c_void_p(clocks.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream)), "probe")
        torch.cuda.synchronize()
        clk = float(clocks.double().mean())
        copies = iters * (nthreads if mode == 1 else 1)
        rows.append({"blocks": blocks, "mode": {1: "ring per thread", 2: "burst on one barrier, same-thread wait", 3: "burst, other-warp wait"}[mode],
                     "issuing_threads": nthreads, "bytes": nbytes, "slots": slots, "clocks_per_copy": round(clk / copies, 1),
                     "bytes_per_clock_per_sm": round(nbytes * copies / clk, 2)})
print(json.dumps(rows, indent=1))
