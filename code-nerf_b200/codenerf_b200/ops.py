"""torch.autograd.Function wrappers over the C ABI of libcnrf.so.

torch is plumbing here: it owns device memory, the current stream and the autograd graph.  Every arithmetic
operation of the hot path happens inside libcnrf.so; a non-CUDA tensor or a missing library raises.
"""
import ctypes as C
import os

import numpy as np
import torch

from . import _lib
from ._lib import FieldShape, FlexPtrs, FlexShape, PtrTable, RenderInputs, RenderShape, check

_f32 = torch.float32
# Grad mode of the CALLER: inside Function.forward autograd is always off, and ctx.needs_input_grad ignores no_grad, so the
# public wrappers record torch.is_grad_enabled() here before .apply (a no_grad render must not write backward images).
_OUTER_GRAD = [True]


def _apply(fn, *args):
    _OUTER_GRAD[0] = torch.is_grad_enabled()
    try:
        return fn.apply(*args)
    finally:
        _OUTER_GRAD[0] = True


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _req_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise _lib.CnrfError("libcnrf operators need CUDA tensors (no CPU fallback on this path)")


def _c(t):
    """fp32 + contiguous view of a tensor (no copy when it already is)."""
    if t.dtype != _f32:
        t = t.to(_f32)
    return t if t.is_contiguous() else t.contiguous()


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _rows(t):
    """(tensor, row stride) for a 2-D operand whose rows may be broadcast (stride 0) -- eval.py:149 passes
    `code.expand(N, -1)`; materialising that copy is exactly what the reference's forward_pass does and we avoid."""
    if t.dtype != _f32:
        t = t.to(_f32)
    if t.dim() == 2 and t.stride(1) == 1 and (t.stride(0) == 0 or t.stride(0) >= t.shape[1]):
        return t, t.stride(0)
    t = t.contiguous()
    return t, t.stride(0)


# ----------------------------------------------------------------------------------------------------------
class _RayBundle(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pose, H, W, focal, cx, cy, select, n, want_ro):
        _req_cuda(pose, select)
        lib = _lib.load()
        pose = _c(pose)
        B = pose.shape[0]
        rd = torch.empty((B * n, 3), dtype=_f32, device=pose.device)
        ro = torch.empty((B * n, 3), dtype=_f32, device=pose.device) if want_ro else None
        check(lib.cnrf_ray_bundle_fwd(_p(pose), B, H, W, focal, cx, cy, _p(select), n, _p(ro), _p(rd), _stream()),
              "cnrf_ray_bundle_fwd")
        ctx.args = (B, H, W, focal, cx, cy, n)
        ctx.select = select
        if want_ro:
            return ro, rd
        dummy = rd.new_empty(0)
        ctx.mark_non_differentiable(dummy)
        return dummy, rd

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g_ro, g_rd):
        B, H, W, focal, cx, cy, n = ctx.args
        lib = _lib.load()
        if g_rd is None:
            g_rd = torch.zeros((B * n, 3), dtype=_f32, device=g_ro.device)
        g_rd = _c(g_rd)
        g_ro = _c(g_ro) if (g_ro is not None and g_ro.numel() == B * n * 3) else None
        g_pose = torch.empty((B, 4, 4), dtype=_f32, device=g_rd.device)
        check(lib.cnrf_ray_bundle_bwd(_p(g_ro), _p(g_rd), B, H, W, focal, cx, cy, _p(ctx.select), n, _p(g_pose),
                                      _stream()), "cnrf_ray_bundle_bwd")
        return g_pose, None, None, None, None, None, None, None, None


def ray_bundle(pose, H, W, focal, cx, cy, select=None, n=None, want_ro=True):
    """rd (and ro) of shape (B*n, 3).  select: int64 CUDA tensor (B, n) of flat pixel indices, or None = all pixels."""
    n = H * W if select is None else n
    ro, rd = _RayBundle.apply(pose, H, W, float(focal), float(cx), float(cy), select, int(n), want_ro)
    return (ro if want_ro else None), rd


# ----------------------------------------------------------------------------------------------------------
class _SampleUniform(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ro, rd, z_table, lower, upper, t_rand):
        _req_cuda(ro, rd, z_table, t_rand)
        lib = _lib.load()
        ro, rd = _c(ro), _c(rd)
        N, S = ro.shape[0], z_table.shape[0]
        pts = torch.empty((N, S, 3), dtype=_f32, device=ro.device)
        z = torch.empty((N, S), dtype=_f32, device=ro.device)
        t_rand = _c(t_rand) if t_rand is not None else None
        check(lib.cnrf_sample_uniform_fwd(_p(ro), _p(rd), N, S, _p(z_table), _p(lower), _p(upper), _p(t_rand), _p(pts),
                                          _p(z), _stream()), "cnrf_sample_uniform_fwd")
        ctx.save_for_backward(z)
        ctx.mark_non_differentiable(z)
        return pts, z

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g_pts, _g_z):
        (z,) = ctx.saved_tensors
        g_ro, g_rd = points_bwd(g_pts, z)
        return g_ro, g_rd, None, None, None, None


def points_bwd(g_pts, z):
    lib = _lib.load()
    g_pts = _c(g_pts)
    N, S = z.shape
    g_ro = torch.empty((N, 3), dtype=_f32, device=z.device)
    g_rd = torch.empty((N, 3), dtype=_f32, device=z.device)
    check(lib.cnrf_points_bwd(_p(g_pts), _p(z), N, S, _p(g_ro), _p(g_rd), _stream()), "cnrf_points_bwd")
    return g_ro, g_rd


def sample_uniform(ro, rd, z_table, lower, upper, t_rand):
    return _SampleUniform.apply(ro, rd, z_table, lower, upper, t_rand)


class _SamplePdf(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ro, rd, weights, z_coarse, num_fine, u, debug):
        _req_cuda(ro, rd, weights, z_coarse, u)
        lib = _lib.load()
        ro, rd, z_coarse = _c(ro), _c(rd), _c(z_coarse)
        weights, w_stride = _rows(weights.detach())
        N, Sc = z_coarse.shape
        S = Sc + num_fine
        pts = torch.empty((N, S, 3), dtype=_f32, device=ro.device)
        z = torch.empty((N, S), dtype=_f32, device=ro.device)
        cdf = torch.empty((N, Sc - 1), dtype=_f32, device=ro.device) if debug else None
        inds = torch.empty((N, num_fine), dtype=torch.int64, device=ro.device) if debug else None
        u = _c(u) if u is not None else None
        check(lib.cnrf_sample_pdf_fwd(_p(ro), _p(rd), _p(weights), w_stride, _p(z_coarse), N, Sc, num_fine, _p(u),
                                      _p(pts), _p(z), _p(cdf), _p(inds), _stream()), "cnrf_sample_pdf_fwd")
        ctx.save_for_backward(z)
        ctx.mark_non_differentiable(z)
        if debug:
            ctx.mark_non_differentiable(cdf, inds)
            return pts, z, cdf, inds
        return pts, z

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g_pts, *_):
        (z,) = ctx.saved_tensors
        g_ro, g_rd = points_bwd(g_pts, z)
        return g_ro, g_rd, None, None, None, None, None


def sample_pdf(ro, rd, weights, z_coarse, num_fine, u, debug=False):
    # the samples carry no gradient back to the weights or depths (point_sampler.py:115 detaches them)
    return _SamplePdf.apply(ro, rd, weights.detach(), z_coarse.detach(), int(num_fine), u, bool(debug))


def invert_cdf(cdf, bins, u):
    _req_cuda(cdf, bins, u)
    lib = _lib.load()
    cdf, bins, u = _c(cdf), _c(bins), _c(u)
    N, Cn = cdf.shape
    K = u.shape[1]
    samples = torch.empty((N, K), dtype=_f32, device=cdf.device)
    inds = torch.empty((N, K), dtype=torch.int64, device=cdf.device)
    check(lib.cnrf_invert_cdf(_p(cdf), _p(bins), _p(u), N, Cn, K, _p(samples), _p(inds), _stream()), "cnrf_invert_cdf")
    return samples, inds


# ----------------------------------------------------------------------------------------------------------
class _PosEmbed(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, freqs, include_input):
        _req_cuda(x, freqs)
        lib = _lib.load()
        x = _c(x)
        M, D = x.shape
        L = freqs.shape[0]
        out = torch.empty((M, D * (int(include_input) + 2 * L)), dtype=_f32, device=x.device)
        check(lib.cnrf_pos_embed_fwd(_p(x), M, D, _p(freqs), L, int(include_input), _p(out), _stream()),
              "cnrf_pos_embed_fwd")
        ctx.save_for_backward(x, freqs)
        ctx.inc = int(include_input)
        return out

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g_out):
        x, freqs = ctx.saved_tensors
        lib = _lib.load()
        g_out = _c(g_out)
        g_x = torch.empty_like(x)
        check(lib.cnrf_pos_embed_bwd(_p(x), _p(g_out), x.shape[0], x.shape[1], _p(freqs), freqs.shape[0], ctx.inc,
                                     _p(g_x), _stream()), "cnrf_pos_embed_bwd")
        return g_x, None, None


def pos_embed(x, freqs, include_input):
    return _PosEmbed.apply(x, freqs, bool(include_input))


# ----------------------------------------------------------------------------------------------------------
class _Composite(torch.autograd.Function):
    @staticmethod
    def forward(ctx, raw, z, rd):
        _req_cuda(raw, z, rd)
        lib = _lib.load()
        raw, z, rd = _c(raw), _c(z), _c(rd)
        N, S = z.shape
        dev = raw.device
        rgb = torch.empty((N, 3), dtype=_f32, device=dev)
        disp = torch.empty((N,), dtype=_f32, device=dev)
        acc = torch.empty((N,), dtype=_f32, device=dev)
        weights = torch.empty((N, S), dtype=_f32, device=dev)
        depth = torch.empty((N,), dtype=_f32, device=dev)
        check(lib.cnrf_composite_fwd(_p(raw), _p(z), _p(rd), N, S, _p(rgb), _p(disp), _p(acc), _p(weights), _p(depth),
                                     _stream()), "cnrf_composite_fwd")
        ctx.save_for_backward(raw, z, rd)
        ctx.set_materialize_grads(False)       # unused outputs arrive as None (the kernel takes NULL), not as zero tensors
        return rgb, disp, acc, weights, depth

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g_rgb, g_disp, g_acc, g_weights, g_depth):
        raw, z, rd = ctx.saved_tensors
        lib = _lib.load()
        N, S = z.shape
        gs = [None if g is None else _c(g) for g in (g_rgb, g_disp, g_acc, g_weights, g_depth)]
        g_raw = torch.empty_like(raw)
        g_z = torch.empty_like(z) if ctx.needs_input_grad[1] else None
        g_rd = torch.empty_like(rd) if ctx.needs_input_grad[2] else None
        check(lib.cnrf_composite_bwd(_p(raw), _p(z), _p(rd), N, S, _p(gs[0]), _p(gs[1]), _p(gs[2]), _p(gs[3]), _p(gs[4]),
                                     _p(g_raw), _p(g_z), _p(g_rd), _stream()), "cnrf_composite_bwd")
        return g_raw, g_z, g_rd


def composite(raw, z, rd):
    """-> (rgb, disp, acc, weights, depth), the tuple order of volume_render (volumetric_render.py:66)."""
    return _Composite.apply(raw, z, rd)


class _MsePair(torch.autograd.Function):
    @staticmethod
    def forward(ctx, rgb_a, rgb_b, target):
        _req_cuda(rgb_a, rgb_b, target)
        lib = _lib.load()
        a, lda = _rows(rgb_a)
        b, ldb = _rows(rgb_b)
        t, ldt = _rows(target)
        N = a.shape[0]
        assert a.shape[-1] >= 3 and b.shape == a.shape and t.shape[0] == N and t.shape[-1] >= 3
        out = torch.empty((3,), dtype=_f32, device=a.device)
        check(lib.cnrf_mse_pair_fwd(_p(a), lda, _p(b), ldb, _p(t), ldt, N, _p(out), _stream()), "cnrf_mse_pair_fwd")
        ctx.save_for_backward(a, b, t)
        ctx.lds = (lda, ldb, ldt)
        ctx.set_materialize_grads(False)
        return out[0], out[1], out[2]

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g0, g1, g2):
        a, b, t = ctx.saved_tensors
        lda, ldb, ldt = ctx.lds
        lib = _lib.load()
        N = a.shape[0]
        gs = [None if g is None else _c(g.to(_f32)) for g in (g0, g1, g2)]
        g_a = torch.empty((N, 3), dtype=_f32, device=a.device) if ctx.needs_input_grad[0] else None
        g_b = torch.empty((N, 3), dtype=_f32, device=a.device) if ctx.needs_input_grad[1] else None
        check(lib.cnrf_mse_pair_bwd(_p(a), lda, _p(b), ldb, _p(t), ldt, N, _p(gs[0]), _p(gs[1]), _p(gs[2]), _p(g_a), _p(g_b),
                                    _stream()), "cnrf_mse_pair_bwd")
        if g_a is not None and a.shape[-1] != 3:
            g_a = torch.nn.functional.pad(g_a, (0, a.shape[-1] - 3))
        if g_b is not None and b.shape[-1] != 3:
            g_b = torch.nn.functional.pad(g_b, (0, b.shape[-1] - 3))
        return g_a, g_b, None


def mse_pair(rgb_coarse, rgb_fine, target):
    """-> (coarse_loss, fine_loss, coarse_loss + fine_loss) of train.py:103-104 as 0-dim tensors: the two
    torch.nn.functional.mse_loss(rgb[..., :3], target[..., :3]) calls and their sum in one launch (one more for both
    gradients).  rgb_* are (N, >=3), target is (N, >=3) and may be a strided view (RGBA rows)."""
    return _MsePair.apply(rgb_coarse, rgb_fine, target)


# ----------------------------------------------------------------------------------------------------------
# CodeNeRF field
# ----------------------------------------------------------------------------------------------------------
def _ptr_table(tensors18):
    t = PtrTable()
    for l in range(_lib.NUM_LAYERS):
        t.weight[l] = tensors18[2 * l].data_ptr()
        t.bias[l] = tensors18[2 * l + 1].data_ptr()
    return t


def _bytes(n, device):
    return torch.empty((max(int(n), 1),), dtype=torch.uint8, device=device)


def _shape(hidden, code, dim_xyz, dim_dir, N, S, pairing, precision, keep):
    return FieldShape(int(hidden), int(code), int(dim_xyz), int(dim_dir), int(N), int(S), int(pairing), int(precision),
                      int(keep))


class _Field(torch.autograd.Function):
    """forward_pass (nerf/__init__.py:94-134): pts (N,S,3), rd (N,3), z_s/z_t (N,C) -> raw (N,S,4)."""

    @staticmethod
    def forward(ctx, cfg, pts, rd, z_s, z_t, freqs_xyz, freqs_dir, *params):
        _req_cuda(pts, rd, z_s, z_t, *params)
        lib = _lib.load()
        hidden, code, dim_xyz, dim_dir, pairing, precision = cfg
        pts, rd = _c(pts), _c(rd)
        z_s, ld_zs = _rows(z_s)
        z_t, ld_zt = _rows(z_t)
        params = tuple(_c(p) for p in params)
        N, S = pts.shape[0], pts.shape[1]
        keep = int(_OUTER_GRAD[0] and any(ctx.needs_input_grad))   # under no_grad (rendering) nothing is saved
        shp = _shape(hidden, code, dim_xyz, dim_dir, N, S, pairing, precision, keep)
        saved = _bytes(lib.cnrf_codenerf_saved_bytes(C.byref(shp)), pts.device)
        ws_bytes = lib.cnrf_field_workspace_bytes(C.byref(shp), 0)
        ws = _bytes(ws_bytes, pts.device)
        raw = torch.empty((N, S, 4), dtype=_f32, device=pts.device)
        tab = _ptr_table(params)
        check(lib.cnrf_field_fwd(C.byref(shp), C.byref(tab), _p(pts), _p(rd), _p(freqs_xyz), freqs_xyz.shape[0],
                                 _p(freqs_dir), freqs_dir.shape[0], _p(z_s), ld_zs, _p(z_t), ld_zt, _p(raw),
                                 _p(saved) if keep else None, saved.numel() if keep else 0, _p(ws), ws.numel(),
                                 _stream()), "cnrf_field_fwd")
        if keep:
            ctx.save_for_backward(pts, rd, z_s, z_t, freqs_xyz, freqs_dir, saved, *params)
            ctx.cfg, ctx.lds = cfg, (ld_zs, ld_zt)
        return raw

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g_raw):
        pts, rd, z_s, z_t, freqs_xyz, freqs_dir, saved, *params = ctx.saved_tensors
        lib = _lib.load()
        hidden, code, dim_xyz, dim_dir, pairing, precision = ctx.cfg
        ld_zs, ld_zt = ctx.lds
        N, S = pts.shape[0], pts.shape[1]
        shp = _shape(hidden, code, dim_xyz, dim_dir, N, S, pairing, precision, 1)
        g_raw = _c(g_raw)
        dev = pts.device
        # a frozen network (eval.py:121-124) needs no weight gradients: the fp32-layout chains then skip those GEMMs
        frozen = precision != _lib.PRECISION_BF16 and not any(ctx.needs_input_grad[7:])
        grads = (None,) * len(params) if frozen else tuple(torch.empty_like(p) for p in params)
        g_zs = torch.empty((N, code), dtype=_f32, device=dev)
        g_zt = torch.empty((N, code), dtype=_f32, device=dev)
        g_pts = torch.empty_like(pts) if ctx.needs_input_grad[1] else None
        g_rd = torch.empty_like(rd) if ctx.needs_input_grad[2] else None
        ws = _bytes(lib.cnrf_field_workspace_bytes(C.byref(shp), 1), dev)
        tab = _ptr_table(params)
        gtab = None if frozen else C.byref(_ptr_table(grads))
        check(lib.cnrf_field_bwd(C.byref(shp), C.byref(tab), _p(pts), _p(rd), _p(freqs_xyz), freqs_xyz.shape[0],
                                 _p(freqs_dir), freqs_dir.shape[0], _p(z_s), ld_zs, _p(z_t), ld_zt, _p(g_raw), _p(saved),
                                 saved.numel(), gtab, _p(g_zs), _p(g_zt), _p(g_pts), _p(g_rd), _p(ws),
                                 ws.numel(), _stream()), "cnrf_field_bwd")
        return (None, g_pts, g_rd, g_zs, g_zt, None, None) + grads


class _FieldIds(torch.autograd.Function):
    """forward_pass with the embedding lookup of train.py:96 folded in: ids (N,) int64, tab_s / tab_t (num_codes, C) are
    the ShapeTextureEmbedding tables.  Rays with the same id as their predecessor share the code-layer work, and the
    backward returns dense table gradients directly (what torch.nn.Embedding's backward accumulates)."""

    @staticmethod
    def forward(ctx, cfg, pts, rd, ids, tab_s, tab_t, freqs_xyz, freqs_dir, *params):
        _req_cuda(pts, rd, ids, tab_s, tab_t, *params)
        lib = _lib.load()
        hidden, code, dim_xyz, dim_dir, pairing, precision = cfg
        pts, rd, ids = _c(pts), _c(rd), ids.contiguous()
        tab_s, ld_s = _rows(tab_s)
        tab_t, ld_t = _rows(tab_t)
        params = tuple(_c(p) for p in params)
        N, S = pts.shape[0], pts.shape[1]
        assert ids.dtype == torch.int64 and ids.shape == (N,), "ids must be (N,) int64"
        assert tab_s.shape[0] == tab_t.shape[0], "shape and texture tables must have the same number of rows"
        keep = int(_OUTER_GRAD[0] and any(ctx.needs_input_grad))   # under no_grad (rendering) nothing is saved
        shp = _shape(hidden, code, dim_xyz, dim_dir, N, S, pairing, precision, keep)
        saved = _bytes(lib.cnrf_codenerf_saved_bytes(C.byref(shp)), pts.device)
        ws = _bytes(lib.cnrf_field_workspace_bytes(C.byref(shp), 0), pts.device)
        raw = torch.empty((N, S, 4), dtype=_f32, device=pts.device)
        tab = _ptr_table(params)
        check(lib.cnrf_field_fwd_ids(C.byref(shp), C.byref(tab), _p(pts), _p(rd), _p(freqs_xyz), freqs_xyz.shape[0],
                                     _p(freqs_dir), freqs_dir.shape[0], _p(ids), tab_s.shape[0], _p(tab_s), ld_s,
                                     _p(tab_t), ld_t, _p(raw), _p(saved) if keep else None, saved.numel() if keep else 0,
                                     _p(ws), ws.numel(), _stream()), "cnrf_field_fwd_ids")
        if keep:
            ctx.save_for_backward(pts, rd, ids, tab_s, tab_t, freqs_xyz, freqs_dir, saved, *params)
            ctx.cfg, ctx.lds = cfg, (ld_s, ld_t)
        return raw

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g_raw):
        pts, rd, ids, tab_s, tab_t, freqs_xyz, freqs_dir, saved, *params = ctx.saved_tensors
        lib = _lib.load()
        hidden, code, dim_xyz, dim_dir, pairing, precision = ctx.cfg
        ld_s, ld_t = ctx.lds
        N, S = pts.shape[0], pts.shape[1]
        shp = _shape(hidden, code, dim_xyz, dim_dir, N, S, pairing, precision, 1)
        g_raw = _c(g_raw)
        dev = pts.device
        grads = tuple(torch.empty_like(p) for p in params)
        g_ts = torch.empty((tab_s.shape[0], code), dtype=_f32, device=dev)
        g_tt = torch.empty((tab_t.shape[0], code), dtype=_f32, device=dev)
        ws = _bytes(lib.cnrf_field_workspace_bytes(C.byref(shp), 1), dev)
        tab, gtab = _ptr_table(params), _ptr_table(grads)
        check(lib.cnrf_field_bwd_ids(C.byref(shp), C.byref(tab), _p(pts), _p(rd), _p(freqs_xyz), freqs_xyz.shape[0],
                                     _p(freqs_dir), freqs_dir.shape[0], _p(ids), tab_s.shape[0], _p(tab_s), ld_s,
                                     _p(tab_t), ld_t, _p(g_raw), _p(saved), saved.numel(), C.byref(gtab), _p(g_ts),
                                     _p(g_tt), _p(ws), ws.numel(), _stream()), "cnrf_field_bwd_ids")
        return (None, None, None, None, g_ts, g_tt, None, None) + grads


POSE_GRAD_PRECISION_DEFAULT = "split"


def codenerf_field(params, pts, rd, z_s, z_t, freqs_xyz, freqs_dir, hidden, code, dim_xyz, dim_dir,
                   pairing=_lib.PAIRING_REFERENCE, precision=_lib.PRECISION_BF16, code_source=None):
    """params: the 18 tensors (weight, bias per layer) in _lib.LAYER_NAMES order.

    code_source = (ids, shape_table, texture_table): z_s / z_t are the rows ids of those tables (ShapeTextureEmbedding's
    output); the bf16 path then shares the code-layer work between rays of equal id (_FieldIds)."""
    pose_grad = (pts.requires_grad or rd.requires_grad) and torch.is_grad_enabled()
    if precision == _lib.PRECISION_BF16 and pose_grad:
        # camera-pose optimisation (eval.py:129-168): the point gradient sums encoding terms weighted by frequencies up
        # to 512, and the bf16 chain's fifth backward stage reproduces the fp32 ray gradients only to ~4 % (cosine 0.999,
        # tests/test_gpu_tc.py::test_tc_pose_gradients_match_fp32).  CODENERF_POSE_GRAD_PRECISION selects the chain such
        # a call runs on: "fp32" CUDA-core FMA (the golden parity mode), "split" the same path with split-bf16
        # tensor-core contractions (~16 mantissa bits), "bf16" the fused chain.
        precision = {"fp32": _lib.PRECISION_FP32, "split": _lib.PRECISION_SPLIT, "bf16": _lib.PRECISION_BF16}[
            os.environ.get("CODENERF_POSE_GRAD_PRECISION", POSE_GRAD_PRECISION_DEFAULT)]
    if pose_grad:
        code_source = None                # the shared-code path carries no ray / point gradients
    cfg = (hidden, code, dim_xyz, dim_dir, pairing, precision)
    if code_source is not None and precision == _lib.PRECISION_BF16:
        ids, tab_s, tab_t = code_source
        if ids.dim() == 1 and ids.shape[0] == pts.shape[0] and ids.dtype == torch.int64:
            return _apply(_FieldIds, cfg, pts, rd, ids, tab_s, tab_t, freqs_xyz, freqs_dir, *params)
    return _apply(_Field, cfg, pts, rd, z_s, z_t, freqs_xyz, freqs_dir, *params)


# ----------------------------------------------------------------------------------------------------------
# The whole pipeline of predict_radiance_and_render (nerf/__init__.py:74-91) as one library call per direction
# ----------------------------------------------------------------------------------------------------------
def _flat_grads(params, dev):
    """One allocation for all parameter gradients, every gradient 16-byte aligned inside it (fc_out's 257-element bias would
    otherwise leave everything behind it on 4-byte boundaries: the weight-gradient kernel's vector reductions and FusedAdamW's
    float4 path both want aligned rows)."""
    offs, total = [], 0
    for p in params:
        offs.append(total)
        total += (p.numel() + 3) & ~3
    flat = torch.empty((total,), dtype=_f32, device=dev)
    return flat, tuple(flat[o:o + p.numel()].view_as(p) for o, p in zip(offs, params))


class RenderSpec(object):
    """Everything about one fused render call that is not a differentiable tensor."""
    __slots__ = ("hidden", "code", "dim_xyz", "dim_dir", "num_coarse", "num_fine", "pairing", "precision", "z_table", "lower",
                 "upper", "freqs_xyz", "freqs_dir", "ids", "want_extras")

    def __init__(self, hidden, code, dim_xyz, dim_dir, num_coarse, num_fine, pairing, precision, z_table, lower, upper,
                 freqs_xyz, freqs_dir, ids=None, want_extras=False):
        self.hidden, self.code, self.dim_xyz, self.dim_dir = int(hidden), int(code), int(dim_xyz), int(dim_dir)
        self.num_coarse, self.num_fine, self.pairing, self.precision = int(num_coarse), int(num_fine), int(pairing), int(precision)
        self.z_table, self.lower, self.upper, self.freqs_xyz, self.freqs_dir = z_table, lower, upper, freqs_xyz, freqs_dir
        self.ids, self.want_extras = ids, bool(want_extras)


def _render_inputs(spec, ro, rd, t_rand, u, code_a, ld_a, code_b, ld_b):
    r = RenderInputs()
    r.ro, r.rd = ro.data_ptr() if ro is not None else None, rd.data_ptr()
    r.z_table, r.lower, r.upper = spec.z_table.data_ptr(), spec.lower.data_ptr(), spec.upper.data_ptr()
    r.t_rand = t_rand.data_ptr() if t_rand is not None else None
    r.u = u.data_ptr() if u is not None else None
    r.freqs_xyz, r.L_xyz = spec.freqs_xyz.data_ptr(), spec.freqs_xyz.shape[0]
    r.freqs_dir, r.L_dir = spec.freqs_dir.data_ptr(), spec.freqs_dir.shape[0]
    r.ids = spec.ids.data_ptr() if spec.ids is not None else None
    r.num_codes = code_a.shape[0] if spec.ids is not None else 0
    r.z_s, r.ld_zs, r.z_t, r.ld_zt = code_a.data_ptr(), ld_a, code_b.data_ptr(), ld_b
    return r


class _Render(torch.autograd.Function):
    """(ro, rd, [t_rand, u], codes, 2 x 18 parameters) -> rgb_coarse, rgb_fine [, extras_fine (3, N)].  code_a / code_b are the
    two embedding tables when spec.ids is set (shared-code mode; dense table gradients come back), per-ray code rows
    otherwise."""

    @staticmethod
    def forward(ctx, spec, ro, rd, t_rand, u, code_a, code_b, *params):
        _req_cuda(ro, rd, t_rand, u, code_a, code_b, *params)
        lib = _lib.load()
        ro, rd = _c(ro), _c(rd)
        t_rand = _c(t_rand) if t_rand is not None else None
        u = _c(u) if u is not None else None
        code_a, ld_a = _rows(code_a)
        code_b, ld_b = _rows(code_b)
        params = tuple(_c(p) for p in params)
        N = ro.shape[0]
        dev = ro.device
        if spec.ids is not None:
            assert spec.ids.dtype == torch.int64 and spec.ids.shape == (N,) and spec.ids.is_contiguous(), "ids must be (N,) int64"
            assert code_a.shape[0] == code_b.shape[0], "shape and texture tables must have the same number of rows"
        keep = int(_OUTER_GRAD[0] and any(ctx.needs_input_grad))
        shp = RenderShape(spec.hidden, spec.code, spec.dim_xyz, spec.dim_dir, N, spec.num_coarse, spec.num_fine, spec.pairing,
                          spec.precision, keep)
        saved = _bytes(lib.cnrf_render_saved_bytes(C.byref(shp)), dev) if keep else None
        ws = _bytes(lib.cnrf_render_workspace_bytes(C.byref(shp), 0), dev)
        rgb = (torch.empty((N, 3), dtype=_f32, device=dev), torch.empty((N, 3), dtype=_f32, device=dev))
        extras = torch.empty((3, N), dtype=_f32, device=dev) if spec.want_extras else None
        tc, tf = _ptr_table(params[:18]), _ptr_table(params[18:])
        inp = _render_inputs(spec, ro, rd, t_rand, u, code_a, ld_a, code_b, ld_b)
        check(lib.cnrf_render_fwd(C.byref(shp), C.byref(tc), C.byref(tf), C.byref(inp), _p(rgb[0]), _p(rgb[1]), _p(extras),
                                  _p(saved), saved.numel() if keep else 0, _p(ws), ws.numel(), _stream()), "cnrf_render_fwd")
        if keep:
            ctx.save_for_backward(rd, code_a, code_b, saved, *params)
            ctx.spec, ctx.lds, ctx.N = spec, (ld_a, ld_b), N
        ctx.set_materialize_grads(False)
        if extras is not None:
            ctx.mark_non_differentiable(extras)
            return rgb[0], rgb[1], extras
        return rgb[0], rgb[1]

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g_c, g_f, *_):
        rd, code_a, code_b, saved, *params = ctx.saved_tensors
        spec, (ld_a, ld_b), N = ctx.spec, ctx.lds, ctx.N
        lib = _lib.load()
        dev = rd.device
        if g_c is None and g_f is None:
            return (None,) * (7 + len(params))
        g_c = _c(g_c) if g_c is not None else None
        g_f = _c(g_f) if g_f is not None else None
        shp = RenderShape(spec.hidden, spec.code, spec.dim_xyz, spec.dim_dir, N, spec.num_coarse, spec.num_fine, spec.pairing,
                          spec.precision, 1)
        flat, grads = _flat_grads(params, dev)                             # all 36 parameter gradients: one allocation
        rows = code_a.shape[0] if spec.ids is not None else N
        g_codes = torch.empty((2, rows, spec.code), dtype=_f32, device=dev)
        ws = _bytes(lib.cnrf_render_workspace_bytes(C.byref(shp), 1), dev)
        tc, tf = _ptr_table(params[:18]), _ptr_table(params[18:])
        gc, gf = _ptr_table(grads[:18]), _ptr_table(grads[18:])
        inp = _render_inputs(spec, None, rd, None, None, code_a, ld_a, code_b, ld_b)
        check(lib.cnrf_render_bwd(C.byref(shp), C.byref(tc), C.byref(tf), C.byref(inp), _p(g_c), _p(g_f), _p(saved), saved.numel(),
                                  C.byref(gc), C.byref(gf), _p(g_codes[0]), _p(g_codes[1]), _p(ws), ws.numel(), _stream()),
              "cnrf_render_bwd")
        return (None, None, None, None, None, g_codes[0], g_codes[1]) + grads


class LossTargetSpec(object):
    """Target pixels of a fused render + loss call: `rows` (N, >= 3) float32 rows already gathered (train.py:77-80), or
    `views` (any shape ending in `channels` uint8, e.g. DeviceSRN.colors_u8) + `pixel` (N,) int64 flat pixel indices
    (view index * H W + selected pixel): the gather and the loader's / 255.0 then happen inside the compositing epilogue."""
    __slots__ = ("rows", "views", "pixel")

    def __init__(self, rows=None, views=None, pixel=None):
        assert (rows is None) != (views is None), "give target rows or uint8 views + pixel indices"
        if views is not None:
            assert pixel is not None and views.dtype == torch.uint8 and pixel.dtype == torch.int64 and views.shape[-1] in (3, 4)
        self.rows, self.views, self.pixel = rows, views, pixel

    def struct(self, N, keepalive):
        t = _lib.LossTarget()
        if self.views is not None:
            v, px = self.views.contiguous(), self.pixel.contiguous()      # uint8 / int64: no dtype conversion
            assert v.is_cuda and px.is_cuda, "views / pixel must be CUDA tensors (there is no CPU fallback)"
            assert px.shape == (N,), "pixel must hold one index per ray"
            keepalive += [v, px]
            t.views, t.pixel, t.channels = v.data_ptr(), px.data_ptr(), v.shape[-1]
        else:
            r, ld = _rows(self.rows)
            assert r.shape[0] == N and r.shape[-1] >= 3 and r.dtype == _f32
            keepalive.append(r)
            t.rows, t.ld_rows = r.data_ptr(), ld
        return t


class _RenderLoss(torch.autograd.Function):
    """_Render + the photometric loss of train.py:103-104 folded into the compositing epilogues (cnrf_render_loss_fwd / _bwd):
    -> coarse_loss, fine_loss, loss (0-dim), rgb_coarse, rgb_fine (not differentiable: their gradient is the loss's)."""

    @staticmethod
    def forward(ctx, spec, target, ro, rd, t_rand, u, code_a, code_b, *params):
        _req_cuda(ro, rd, t_rand, u, code_a, code_b, *params)
        lib = _lib.load()
        ro, rd = _c(ro), _c(rd)
        t_rand = _c(t_rand) if t_rand is not None else None
        u = _c(u) if u is not None else None
        code_a, ld_a = _rows(code_a)
        code_b, ld_b = _rows(code_b)
        params = tuple(_c(p) for p in params)
        N = ro.shape[0]
        dev = ro.device
        if spec.ids is not None:
            assert spec.ids.dtype == torch.int64 and spec.ids.shape == (N,) and spec.ids.is_contiguous(), "ids must be (N,) int64"
            assert code_a.shape[0] == code_b.shape[0], "shape and texture tables must have the same number of rows"
        keep = int(_OUTER_GRAD[0] and any(ctx.needs_input_grad))
        shp = RenderShape(spec.hidden, spec.code, spec.dim_xyz, spec.dim_dir, N, spec.num_coarse, spec.num_fine, spec.pairing,
                          spec.precision, keep)
        saved = _bytes(lib.cnrf_render_saved_bytes(C.byref(shp)), dev) if keep else None
        ws = _bytes(lib.cnrf_render_workspace_bytes(C.byref(shp), 0), dev)
        rgb = torch.empty((2, N, 3), dtype=_f32, device=dev)
        loss3 = torch.empty((3,), dtype=_f32, device=dev)
        tc, tf = _ptr_table(params[:18]), _ptr_table(params[18:])
        inp = _render_inputs(spec, ro, rd, t_rand, u, code_a, ld_a, code_b, ld_b)
        alive = []
        tgt = target.struct(N, alive)
        check(lib.cnrf_render_loss_fwd(C.byref(shp), C.byref(tc), C.byref(tf), C.byref(inp), C.byref(tgt), _p(loss3), _p(rgb[0]),
                                       _p(rgb[1]), _p(saved), saved.numel() if keep else 0, _p(ws), ws.numel(), _stream()),
              "cnrf_render_loss_fwd")
        if keep:
            ctx.save_for_backward(rd, code_a, code_b, saved, rgb, *alive, *params)
            ctx.spec, ctx.lds, ctx.N, ctx.n_alive, ctx.target = spec, (ld_a, ld_b), N, len(alive), target
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(rgb)
        return loss3[0], loss3[1], loss3[2], rgb

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g0, g1, g2, _g_rgb):
        rd, code_a, code_b, saved, rgb, *rest = ctx.saved_tensors
        alive, params = rest[:ctx.n_alive], rest[ctx.n_alive:]
        spec, (ld_a, ld_b), N = ctx.spec, ctx.lds, ctx.N
        lib = _lib.load()
        dev = rd.device
        if g0 is None and g1 is None and g2 is None:
            return (None,) * (8 + len(params))
        gs = [None if g is None else _c(g.to(_f32)) for g in (g0, g1, g2)]
        shp = RenderShape(spec.hidden, spec.code, spec.dim_xyz, spec.dim_dir, N, spec.num_coarse, spec.num_fine, spec.pairing,
                          spec.precision, 1)
        flat, grads = _flat_grads(params, dev)                             # all 36 parameter gradients: one allocation
        rows = code_a.shape[0] if spec.ids is not None else N
        g_codes = torch.empty((2, rows, spec.code), dtype=_f32, device=dev)
        ws = _bytes(lib.cnrf_render_workspace_bytes(C.byref(shp), 1), dev)
        tc, tf = _ptr_table(params[:18]), _ptr_table(params[18:])
        gc, gf = _ptr_table(grads[:18]), _ptr_table(grads[18:])
        inp = _render_inputs(spec, None, rd, None, None, code_a, ld_a, code_b, ld_b)
        tgt = ctx.target.struct(N, [])
        check(lib.cnrf_render_loss_bwd(C.byref(shp), C.byref(tc), C.byref(tf), C.byref(inp), C.byref(tgt), _p(rgb[0]), _p(rgb[1]),
                                       _p(gs[0]), _p(gs[1]), _p(gs[2]), _p(saved), saved.numel(), C.byref(gc), C.byref(gf),
                                       _p(g_codes[0]), _p(g_codes[1]), _p(ws), ws.numel(), _stream()), "cnrf_render_loss_bwd")
        return (None, None, None, None, None, None, g_codes[0], g_codes[1]) + grads


def render_loss(spec, target, ro, rd, t_rand, u, code_a, code_b, params_coarse, params_fine):
    """predict_radiance_and_render + train.py:103-104's photometric loss as one library call per direction:
    -> (coarse_loss, fine_loss, loss, rgb (2, N, 3) [coarse, fine; detached])."""
    return _apply(_RenderLoss, spec, target, ro, rd, t_rand, u, code_a, code_b, *params_coarse, *params_fine)


def render(spec, ro, rd, t_rand, u, code_a, code_b, params_coarse, params_fine):
    """predict_radiance_and_render as one library call: -> (rgb_coarse, rgb_fine[, extras_fine])."""
    return _apply(_Render, spec, ro, rd, t_rand, u, code_a, code_b, *params_coarse, *params_fine)


class _Model(torch.autograd.Function):
    """CodeNeRFModel.forward(z_s, z_t, x) as written (models/model.py:160-194): per-sample codes, pre-embedded x."""

    @staticmethod
    def forward(ctx, cfg, z_s, z_t, x, *params):
        _req_cuda(z_s, z_t, x, *params)
        lib = _lib.load()
        hidden, code, dim_xyz, dim_dir, precision = cfg
        x = _c(x)
        z_s, ld_zs = _rows(z_s)
        z_t, ld_zt = _rows(z_t)
        params = tuple(_c(p) for p in params)
        M = x.shape[0]
        keep = int(_OUTER_GRAD[0] and any(ctx.needs_input_grad))   # under no_grad (rendering) nothing is saved
        shp = _shape(hidden, code, dim_xyz, dim_dir, M, 1, _lib.PAIRING_CORRECT, precision, keep)
        saved = _bytes(lib.cnrf_codenerf_saved_bytes(C.byref(shp)), x.device)
        ws = _bytes(lib.cnrf_codenerf_workspace_bytes(C.byref(shp), 0), x.device)
        raw = torch.empty((M, 4), dtype=_f32, device=x.device)
        ld = x.shape[1]
        x_dir = C.c_void_p(x.data_ptr() + 4 * dim_xyz)
        tab = _ptr_table(params)
        check(lib.cnrf_codenerf_fwd(C.byref(shp), C.byref(tab), _p(x), ld, x_dir, ld, _p(z_s), ld_zs, _p(z_t), ld_zt,
                                    _p(raw), _p(saved) if keep else None, saved.numel() if keep else 0, _p(ws),
                                    ws.numel(), _stream()), "cnrf_codenerf_fwd")
        if keep:
            ctx.save_for_backward(z_s, z_t, x, saved, *params)
            ctx.cfg, ctx.lds = cfg, (ld_zs, ld_zt)
        return raw

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g_raw):
        z_s, z_t, x, saved, *params = ctx.saved_tensors
        lib = _lib.load()
        hidden, code, dim_xyz, dim_dir, precision = ctx.cfg
        ld_zs, ld_zt = ctx.lds
        M, ld = x.shape
        shp = _shape(hidden, code, dim_xyz, dim_dir, M, 1, _lib.PAIRING_CORRECT, precision, 1)
        g_raw = _c(g_raw)
        dev = x.device
        grads = tuple(torch.empty_like(p) for p in params)
        g_zs = torch.empty((M, code), dtype=_f32, device=dev)
        g_zt = torch.empty((M, code), dtype=_f32, device=dev)
        need_x = ctx.needs_input_grad[3]
        g_xyz = torch.empty((M, dim_xyz), dtype=_f32, device=dev) if need_x else None
        g_dir = torch.empty((M, dim_dir), dtype=_f32, device=dev) if need_x else None
        ws = _bytes(lib.cnrf_codenerf_workspace_bytes(C.byref(shp), 1), dev)
        x_dir = C.c_void_p(x.data_ptr() + 4 * dim_xyz)
        tab, gtab = _ptr_table(params), _ptr_table(grads)
        check(lib.cnrf_codenerf_bwd(C.byref(shp), C.byref(tab), _p(x), ld, x_dir, ld, _p(z_s), ld_zs, _p(z_t), ld_zt,
                                    _p(g_raw), _p(saved), saved.numel(), C.byref(gtab), _p(g_zs), _p(g_zt), _p(g_xyz),
                                    _p(g_dir), _p(ws), ws.numel(), _stream()), "cnrf_codenerf_bwd")
        g_x = torch.cat((g_xyz, g_dir), dim=-1) if need_x else None
        return (None, g_zs, g_zt, g_x) + grads


def codenerf_model(params, z_s, z_t, x, hidden, code, dim_xyz, dim_dir, precision=_lib.PRECISION_FP32):
    cfg = (hidden, code, dim_xyz, dim_dir, precision)
    return _apply(_Model, cfg, z_s, z_t, x, *params)


# ----------------------------------------------------------------------------------------------------------
# Vanilla NeRF field (FlexibleNeRFModel, models/model.py:5-76)
# ----------------------------------------------------------------------------------------------------------
def _flex_ptrs(tensors, num_xyz, viewdirs=True):
    """tensors: (w,b) pairs in state_dict order -- layer1, layers_xyz[0..], fc_feat, layers_dir[0], fc_alpha, fc_rgb; or, for
    the use_viewdirs=False network (dim_dir == 0 at the C boundary), layer1, layers_xyz[0..], fc_out (in the w_rgb slot)."""
    t = FlexPtrs()
    it = iter(tensors)
    t.w_layer1, t.b_layer1 = next(it).data_ptr(), next(it).data_ptr()
    for i in range(num_xyz):
        t.w_xyz[i], t.b_xyz[i] = next(it).data_ptr(), next(it).data_ptr()
    if not viewdirs:
        t.w_rgb, t.b_rgb = next(it).data_ptr(), next(it).data_ptr()
        return t
    t.w_feat, t.b_feat = next(it).data_ptr(), next(it).data_ptr()
    t.w_dir, t.b_dir = next(it).data_ptr(), next(it).data_ptr()
    t.w_alpha, t.b_alpha = next(it).data_ptr(), next(it).data_ptr()
    t.w_rgb, t.b_rgb = next(it).data_ptr(), next(it).data_ptr()
    return t


class _FlexNerf(torch.autograd.Function):
    """x_xyz (M, dim_xyz) embedded points, x_dir (N, dim_dir) embedded per-ray directions, M = N*S -> raw (M, 4)."""

    @staticmethod
    def forward(ctx, cfg, x_xyz, x_dir, *params):
        _req_cuda(x_xyz, x_dir, *params)
        lib = _lib.load()
        num_layers, hidden, dim_xyz, dim_dir, skip_mask, S, pairing, precision = cfg
        x_xyz, ld_xyz = _rows(x_xyz)
        if ld_xyz == 0:
            x_xyz, ld_xyz = x_xyz.contiguous(), x_xyz.shape[1]
        if dim_dir == 0:         # use_viewdirs=False: x_dir only carries the ray count, (N, 0)
            x_dir, ld_dir = x_dir.new_empty((x_dir.shape[0], 0)), 0
        else:
            x_dir, ld_dir = _rows(x_dir)
            if ld_dir == 0:
                x_dir, ld_dir = x_dir.contiguous(), x_dir.shape[1]
        params = tuple(_c(p) for p in params)
        M, N = x_xyz.shape[0], x_dir.shape[0]
        assert M == N * S, "FlexibleNeRF field: rows of x_xyz must be rays x samples"
        assert len(params) == 2 * (num_layers + (4 if dim_dir else 1)), "FlexibleNeRF field: parameter count"
        keep = int(_OUTER_GRAD[0] and any(ctx.needs_input_grad))   # under no_grad (rendering) nothing is saved
        shp = FlexShape(num_layers, hidden, dim_xyz, dim_dir, skip_mask, N, S, pairing, precision, keep)
        dev = x_xyz.device
        saved = _bytes(lib.cnrf_flexnerf_saved_bytes(C.byref(shp)), dev)
        ws = _bytes(lib.cnrf_flexnerf_workspace_bytes(C.byref(shp), 0), dev)
        raw = torch.empty((M, 4), dtype=_f32, device=dev)
        tab = _flex_ptrs(params, num_layers - 1, dim_dir > 0)
        check(lib.cnrf_flexnerf_fwd(C.byref(shp), C.byref(tab), _p(x_xyz), ld_xyz, _p(x_dir) if dim_dir else None, ld_dir, _p(raw),
                                    _p(saved) if keep else None, saved.numel() if keep else 0, _p(ws), ws.numel(), _stream()),
              "cnrf_flexnerf_fwd")
        if keep:
            ctx.save_for_backward(x_xyz, x_dir, saved, *params)
            ctx.cfg, ctx.lds = cfg, (ld_xyz, ld_dir)
        return raw

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g_raw):
        x_xyz, x_dir, saved, *params = ctx.saved_tensors
        lib = _lib.load()
        num_layers, hidden, dim_xyz, dim_dir, skip_mask, S, pairing, precision = ctx.cfg
        ld_xyz, ld_dir = ctx.lds
        M, N = x_xyz.shape[0], x_dir.shape[0]
        shp = FlexShape(num_layers, hidden, dim_xyz, dim_dir, skip_mask, N, S, pairing, precision, 1)
        dev = x_xyz.device
        g_raw = _c(g_raw)
        grads = tuple(torch.empty_like(p) for p in params)
        g_xyz = torch.empty((M, dim_xyz), dtype=_f32, device=dev) if ctx.needs_input_grad[1] else None
        g_dir = torch.empty((N, dim_dir), dtype=_f32, device=dev) if ctx.needs_input_grad[2] and dim_dir else None
        ws = _bytes(lib.cnrf_flexnerf_workspace_bytes(C.byref(shp), 1), dev)
        tab, gtab = _flex_ptrs(params, num_layers - 1, dim_dir > 0), _flex_ptrs(grads, num_layers - 1, dim_dir > 0)
        check(lib.cnrf_flexnerf_bwd(C.byref(shp), C.byref(tab), _p(x_xyz), ld_xyz, _p(x_dir) if dim_dir else None, ld_dir, _p(g_raw), _p(saved),
                                    saved.numel(), C.byref(gtab), _p(g_xyz), _p(g_dir), _p(ws), ws.numel(), _stream()),
              "cnrf_flexnerf_bwd")
        return (None, g_xyz, g_dir) + grads


def flexnerf_fused_inference(num_layers, hidden, dim_xyz, dim_dir, skip_mask, precision):
    """True where cnrf_flexnerf_fwd_pts applies: no gradients wanted and the fused bf16 chain takes this network."""
    if torch.is_grad_enabled() or dim_dir <= 0:
        return False
    shp = FlexShape(int(num_layers), int(hidden), int(dim_xyz), int(dim_dir), int(skip_mask), 1, 1, _lib.PAIRING_REFERENCE,
                    int(precision), 0)
    return bool(_lib.load().cnrf_flexnerf_fused_inference(C.byref(shp)))


def flexnerf_pts(params, pts, freqs_xyz, include_input, x_dir, num_layers, hidden, dim_xyz, dim_dir, skip_mask, num_samples,
                 pairing=_lib.PAIRING_REFERENCE, precision=_lib.PRECISION_BF16):
    """Inference from sample positions pts (M, 3): gamma(x) is produced inside the fused chain (cnrf_flexnerf_fwd_pts).
    Call only where flexnerf_fused_inference(...) is True."""
    _req_cuda(pts, x_dir, freqs_xyz, *params)
    assert not torch.is_grad_enabled(), "flexnerf_pts is the no-grad inference path"
    lib = _lib.load()
    pts, ld_pts = _rows(pts)
    if ld_pts == 0:
        pts, ld_pts = pts.contiguous(), pts.shape[1]
    x_dir, ld_dir = _rows(x_dir)
    if ld_dir == 0:
        x_dir, ld_dir = x_dir.contiguous(), x_dir.shape[1]
    params = tuple(_c(p.detach()) for p in params)
    freqs_xyz = _c(freqs_xyz)
    M, N = pts.shape[0], x_dir.shape[0]
    assert M == N * num_samples and pts.shape[1] == 3, "FlexibleNeRF field: pts must be (rays x samples, 3)"
    shp = FlexShape(int(num_layers), int(hidden), int(dim_xyz), int(dim_dir), int(skip_mask), N, int(num_samples), int(pairing),
                    int(precision), 0)
    dev = pts.device
    ws = _bytes(lib.cnrf_flexnerf_workspace_bytes(C.byref(shp), 0), dev)
    raw = torch.empty((M, 4), dtype=_f32, device=dev)
    tab = _flex_ptrs(params, num_layers - 1, True)
    check(lib.cnrf_flexnerf_fwd_pts(C.byref(shp), C.byref(tab), _p(pts), ld_pts, _p(freqs_xyz), freqs_xyz.numel(),
                                    int(bool(include_input)), _p(x_dir), ld_dir, _p(raw), _p(ws), ws.numel(), _stream()),
          "cnrf_flexnerf_fwd_pts")
    return raw


def flexnerf(params, x_xyz, x_dir, num_layers, hidden, dim_xyz, dim_dir, skip_mask, num_samples,
             pairing=_lib.PAIRING_REFERENCE, precision=_lib.PRECISION_FP32):
    cfg = (int(num_layers), int(hidden), int(dim_xyz), int(dim_dir), int(skip_mask), int(num_samples), int(pairing),
           int(precision))
    return _apply(_FlexNerf, cfg, x_xyz, x_dir, *params)
