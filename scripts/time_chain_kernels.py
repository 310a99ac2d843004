"""Mean duration of the three tensor-core kernels (fine-pass launches: >= 4096 x 160 rows) inside the C2 chunk step,
CUDA events on the launching stream (csrc/ktimer.cu), plus the whole chunk step.

    python scripts/time_chain_kernels.py [reps=20]
"""
import os, sys, runpy
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "code-nerf_b200")); sys.path.insert(0, ROOT)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
sys.argv = [sys.argv[0], "4096", "3", "bf16"]
g = runpy.run_path(os.path.join(ROOT, "scripts", "profile_chunk.py"))     # builds the models and warms up
step, lib = g["step"], g["_lib"].load()
for fine in (True, False):
    rows = 4096 * 160 if fine else 1
    for which in range(3): lib.cnrf_debug_kernel_ms(which, rows, None, None)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): step()
    e1.record()
    torch.cuda.synchronize()
    ms = [lib.cnrf_debug_kernel_ms(which, 0, None, None) * 1e3 for which in range(3)]
    tag = "fine-pass launches" if fine else "all launches (coarse + fine mean)"
    print(f"{tag}: fwd {ms[0]:.1f} us  bwd {ms[1]:.1f} us  dW {ms[2]:.1f} us   chunk step {e0.elapsed_time(e1) / reps:.3f} ms")
