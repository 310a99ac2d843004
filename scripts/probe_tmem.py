"""TMEM read throughput of one SM by tcgen05.ld shape / warps / loads in flight (csrc/tc_probe.cu).
    python scripts/probe_tmem.py > gpurun_out/tmem_probe.json"""
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
rows = []
for blocks in (1, 148):
    for warps in (4, 8, 16):
        for mode in (0, 1, 2):
            iters = 64
            clocks = torch.zeros(blocks, dtype=torch.int64, device=dev)
            sink = torch.zeros(blocks * warps * 32, dtype=torch.int32, device=dev)
            for _ in range(2):
                _lib.check(lib.cnrf_debug_tmem_ld(blocks, warps, iters, mode, C.c_void_p(clocks.data_ptr()), C.c_void_p(sink.data_ptr()),
                                                  C.c_void_p(torch.cuda.current_stream().cuda_stream)), "cnrf_debug_tmem_ld")
            torch.cuda.synchronize()
            clk = float(clocks.double().mean())
            rows.append({"blocks": blocks, "warps": warps, "mode": ["x32 serial", "x32 two in flight", "x64 serial"][mode],
                         "clocks": clk, "bytes_per_clock_per_sm": warps * iters * 65536 / clk})
print(json.dumps(rows, indent=1))
