"""Standalone CUDA-event timing of the HBM-bound stages (ray bundle, stratified / hierarchical sampling, compositing
forward / backward) against the measured HBM peak, with the algorithmic bytes per ray of SURVEY section 8(d).

    python scripts/time_hbm_kernels.py [rays=524288] [reps=20] [out.json]

Working sets exceed the 126 MB L2 at the default ray count (compositing fine pass: 3.2 KB/ray -> 1.7 GB), so every
repetition streams from HBM; no explicit flush.  Calls go through the C ABI directly (caller-owned buffers), so the
timed region holds exactly one kernel launch per repetition.
"""
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "code-nerf_b200")); sys.path.insert(0, ROOT)
from codenerf_b200 import _lib  # noqa: E402
from codenerf_b200.ops import _p, _stream  # noqa: E402



def hbm_peak():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        return 6457.4, "fallback 6457.4 GB/s (earlier driver measurement on this pool)"


def measure(N=524288, reps=20, dev=None, verbose=True):
    """-> list of {kernel, rays, bytes_per_ray, ms, GBps, peak_GBps, frac}; one kernel launch per timed repetition."""
    dev = dev or torch.device("cuda", 0)
    lib = _lib.load()
    peak, _ = hbm_peak()
    torch.manual_seed(0)
    f32 = dict(dtype=torch.float32, device=dev)


    def timed(fn):
        for _ in range(3):
            rc = fn()
            assert rc == 0, f"libcnrf returned {rc}"
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps


    rows = []


    def report(name, bytes_per_ray, ms, n=None):
        n = N if n is None else n
        gbs = bytes_per_ray * n / (ms * 1e-3) / 1e9
        rows.append({"kernel": name, "rays": n, "bytes_per_ray": bytes_per_ray, "ms": ms, "GBps": gbs, "peak_GBps": peak,
                     "frac": gbs / peak})
        if verbose:
            print(f"{name:34s} {n:8d} rays  {bytes_per_ray:6d} B/ray  {ms * 1e3:8.1f} us  {gbs:7.0f} GB/s  {gbs / peak:5.2f} of peak")


    ro = torch.randn(N, 3, **f32) * 0.01 + torch.tensor([0.0, 0.0, 1.3], device=dev)
    rd = torch.nn.functional.normalize(torch.randn(N, 3, **f32) * 0.3 + torch.tensor([0.0, 0.0, -1.0], device=dev), dim=-1)

    # ---- compositing (a11) ----------------------------------------------------------------------------------------
    for S, label in ((32, "coarse"), (160, "fine")):
        raw = torch.randn(N, S, 4, **f32)
        z = torch.sort(torch.rand(N, S, **f32) + 0.8, dim=-1).values
        rgb, disp, acc, depth = torch.empty(N, 3, **f32), torch.empty(N, **f32), torch.empty(N, **f32), torch.empty(N, **f32)
        w = torch.empty(N, S, **f32)
        ms = timed(lambda: lib.cnrf_composite_fwd(_p(raw), _p(z), _p(rd), N, S, _p(rgb), _p(disp), _p(acc), _p(w), _p(depth),
                                                   _stream()))
        report(f"composite_fwd S={S} ({label})", 24 * S + 36, ms)
        g_rgb, g_raw = torch.randn(N, 3, **f32), torch.empty(N, S, 4, **f32)
        ms = timed(lambda: lib.cnrf_composite_bwd(_p(raw), _p(z), _p(rd), N, S, _p(g_rgb), None, None, None, None, _p(g_raw),
                                                   None, None, _stream()))
        report(f"composite_bwd S={S} (rgb grad only)", 36 * S + 24, ms)
        g_w = torch.randn(N, S, **f32)
        g_rd = torch.empty(N, 3, **f32)
        ms = timed(lambda: lib.cnrf_composite_bwd(_p(raw), _p(z), _p(rd), N, S, _p(g_rgb), None, None, _p(g_w), None, _p(g_raw),
                                                   None, _p(g_rd), _stream()))
        report(f"composite_bwd S={S} (+weights, d rd)", 40 * S + 36, ms)
        del raw, z, w, g_raw, g_w

    # ---- stratified sampling (a5), offsets supplied ---------------------------------------------------------------
    Sc, Nf = 32, 128
    t = torch.linspace(0, 1, Sc, **f32)
    z_tab = 1.0 / (1.0 / 0.8 * (1.0 - t) + 1.0 / 1.8 * t)
    mids = 0.5 * (z_tab[1:] + z_tab[:-1])
    upper, lower = torch.cat([mids, z_tab[-1:]]), torch.cat([z_tab[:1], mids])
    t_rand = torch.rand(N, Sc, **f32)
    pts, zc = torch.empty(N, Sc, 3, **f32), torch.empty(N, Sc, **f32)
    ms = timed(lambda: lib.cnrf_sample_uniform_fwd(_p(ro), _p(rd), N, Sc, _p(z_tab), _p(lower), _p(upper), _p(t_rand), _p(pts),
                                                    _p(zc), _stream()))
    report("sample_uniform S_c=32", 20 * Sc + 24, ms)

    # ---- hierarchical sampling (a6), offsets supplied --------------------------------------------------------------
    wts = torch.rand(N, Sc - 2, **f32)
    u = torch.rand(N, Nf, **f32)
    pts_f, zf = torch.empty(N, Sc + Nf, 3, **f32), torch.empty(N, Sc + Nf, **f32)
    ms = timed(lambda: lib.cnrf_sample_pdf_fwd(_p(ro), _p(rd), _p(wts), Sc - 2, _p(zc), N, Sc, Nf, _p(u), _p(pts_f), _p(zf),
                                                None, None, _stream()))
    report("sample_pdf S_c=32 N_f=128", 4 * (Sc - 2) + 4 * Sc + 4 * Nf + 24 + 16 * (Sc + Nf), ms)

    # ---- points backward (pose optimisation) ------------------------------------------------------------------------
    g_pts = torch.randn(N, Sc + Nf, 3, **f32)
    g_ro, g_rd2 = torch.empty(N, 3, **f32), torch.empty(N, 3, **f32)
    ms = timed(lambda: lib.cnrf_points_bwd(_p(g_pts), _p(zf), N, Sc + Nf, _p(g_ro), _p(g_rd2), _stream()))
    report("points_bwd S=160", 16 * (Sc + Nf) + 24, ms)

    # ---- ray bundle (a2), full images --------------------------------------------------------------------------------
    H = W = 128
    B = max(1, 16 * N // (H * W))        # 16x the rays: a 24 B/ray kernel needs > 100 MB to leave launch latency behind
    pose = torch.eye(4, **f32).expand(B, 4, 4).contiguous()
    ro_b, rd_b = torch.empty(B * H * W, 3, **f32), torch.empty(B * H * W, 3, **f32)
    ms = timed(lambda: lib.cnrf_ray_bundle_fwd(_p(pose), B, H, W, 131.25, 64.0, 64.0, None, H * W, _p(ro_b), _p(rd_b), _stream()))
    report("ray_bundle 128x128 images", 24, ms, n=B * H * W)
    # what a write-only stream of that size reaches on this GPU (the peak in the last column is a COPY figure: read + write):
    # cudaMemsetAsync over the same two buffers (24 B/ray as well)
    def _memset():
        ro_b.zero_(); rd_b.zero_()
        return 0
    ms = timed(_memset)
    report("reference: memset of the same 24 B/ray (write-only ceiling)", 24, ms, n=B * H * W)

    return rows


if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 524288
    r = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    out_path = sys.argv[3] if len(sys.argv) > 3 else None
    res = measure(n, r)
    if out_path:
        pk, src = hbm_peak()
        json.dump({"rays": n, "reps": r, "peak_GBps": pk, "peak_source": src, "kernels": res}, open(out_path, "w"), indent=1)
