"""Per-SM ingest rate of cp.async.bulk (csrc/tc_probe.cu: cnrf_debug_bulk_ld) by copy size, copies in flight, CTAs pulling at
once and source (2 MB walked by every CTA from the same start = the weight-image stream of the chain kernels, L2-resident;
1 GB with per-CTA offsets = the operand stream of the weight-gradient kernel, from HBM).
    python scripts/probe_bulk.py > gpurun_out/bulk_probe.json"""
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "code-nerf_b200"))
from codenerf_b200 import _lib  # noqa: E402

lib = _lib.load()
dev = torch.device("cuda", 0)
props = torch.cuda.get_device_properties(dev)
ghz = None
small = torch.zeros(2 * 1024 * 1024, dtype=torch.uint8, device=dev)
big = torch.zeros(1024 * 1024 * 1024, dtype=torch.uint8, device=dev)
rows = []
st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)
for src_name, src, per_block in (("2MB shared (L2)", small, 0), ("1GB per-CTA offsets (HBM)", big, 1)):
    for blocks in (1, 16, 74, 148):
        for nbytes, slots in ((16384, 1), (16384, 2), (16384, 5), (16384, 10), (8192, 10), (8192, 20 if False else 16), (32768, 5), (4096, 16)):
            if slots * nbytes > 200 * 1024:
                continue
            iters = 2000 if src is small else 600
            clocks = torch.zeros(blocks, dtype=torch.int64, device=dev)
            for _ in range(2):
                _lib.check(lib.cnrf_debug_bulk_ld(blocks, slots, nbytes, iters, C.c_void_p(src.data_ptr()), src.numel(), per_block,
                                                  C.c_void_p(clocks.data_ptr()), st()), "cnrf_debug_bulk_ld")
            torch.cuda.synchronize()
            clk = float(clocks.double().mean())
            rows.append({"source": src_name, "blocks": blocks, "bytes": nbytes, "in_flight": slots,
                         "bytes_per_clock_per_sm": round(nbytes * iters / clk, 2), "clocks_per_copy": round(clk / iters, 1)})
print(json.dumps(rows, indent=1))
