"""tcgen05.mma rate of one SM from static smem operands (N = 128 vs 256), alone and beside bulk copies into smem
(csrc/tc_probe.cu: cnrf_debug_umma_rate).   python scripts/probe_umma_rate.py > gpurun_out/umma_rate_probe.json"""
import ctypes as C, json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "code-nerf_b200"))
from codenerf_b200 import _lib  # noqa: E402
lib = _lib.load()
dev = torch.device("cuda", 0)
src = torch.zeros(2 * 1024 * 1024, dtype=torch.uint8, device=dev)
rows = []
chunks = 4096
for blocks in (1, 148):
    for n_cols in (128, 256):
        for nprod, nbytes in ((0, 16384), (1, 16384), (2, 16384), (3, 16384), (4, 8192)):
            out = torch.zeros(2 * blocks, dtype=torch.int64, device=dev)
            for _ in range(2):
                _lib.check(lib.cnrf_debug_umma_rate(blocks, n_cols, chunks, nprod, nbytes, C.c_void_p(src.data_ptr()), src.numel(),
                                                    C.c_void_p(out.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream)), "probe")
            torch.cuda.synchronize()
            o = out.view(blocks, 2).double()
            clk, copies = float(o[:, 0].mean()), float(o[:, 1].mean())
            rows.append({"blocks": blocks, "mma_N": n_cols, "copy_threads": nprod, "clk_per_K64xN256_chunk": round(clk / chunks, 1),
                         "of_ideal_512": round(512.0 * chunks / clk, 3), "copy_B_per_clk_all_threads": round(copies * nprod * nbytes / clk, 1)})
print(json.dumps(rows, indent=1))
