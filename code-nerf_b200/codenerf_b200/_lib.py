"""ctypes binding of libcnrf.so (include/cnrf.h).  No CPU fallback: if the library is missing, or a call
returns non-zero, this raises."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "libcnrf.so")

NUM_LAYERS = 9
LAYER_NAMES = ("layer_xyz1", "layer_xyz2", "fc_out", "shape_code_layer1", "shape_code_layer2",
               "texture_code_layer1", "layer_dir1", "layer_dir2", "fc_rgb")
PRECISION_FP32, PRECISION_BF16, PRECISION_SPLIT = 0, 1, 2
PAIRING_REFERENCE, PAIRING_CORRECT = 0, 1

c_f32p = C.c_void_p   # device pointers travel as integers
c_i64 = C.c_int64
c_i32 = C.c_int32
c_sz = C.c_size_t
c_f = C.c_float


class FieldShape(C.Structure):
    _fields_ = [("hidden", c_i32), ("code", c_i32), ("dim_xyz", c_i32), ("dim_dir", c_i32), ("num_rays", c_i64),
                ("num_samples", c_i32), ("pairing", c_i32), ("precision", c_i32), ("keep", c_i32)]


class PtrTable(C.Structure):   # cnrf_codenerf_params and cnrf_codenerf_grads share this layout
    _fields_ = [("weight", C.c_void_p * NUM_LAYERS), ("bias", C.c_void_p * NUM_LAYERS)]


FLEX_MAX_XYZ_LAYERS = 15


class FlexShape(C.Structure):   # cnrf_flex_shape
    _fields_ = [("num_layers", c_i32), ("hidden", c_i32), ("dim_xyz", c_i32), ("dim_dir", c_i32), ("skip_mask", C.c_uint32),
                ("num_rays", c_i64), ("num_samples", c_i32), ("pairing", c_i32), ("precision", c_i32), ("keep", c_i32)]


class FlexPtrs(C.Structure):    # cnrf_flex_params and cnrf_flex_grads share this layout
    _fields_ = [("w_layer1", C.c_void_p), ("b_layer1", C.c_void_p), ("w_xyz", C.c_void_p * FLEX_MAX_XYZ_LAYERS),
                ("b_xyz", C.c_void_p * FLEX_MAX_XYZ_LAYERS), ("w_feat", C.c_void_p), ("b_feat", C.c_void_p),
                ("w_dir", C.c_void_p), ("b_dir", C.c_void_p), ("w_alpha", C.c_void_p), ("b_alpha", C.c_void_p),
                ("w_rgb", C.c_void_p), ("b_rgb", C.c_void_p)]

class RenderShape(C.Structure):     # cnrf_render_shape
    _fields_ = [("hidden", c_i32), ("code", c_i32), ("dim_xyz", c_i32), ("dim_dir", c_i32), ("num_rays", c_i64),
                ("num_coarse", c_i32), ("num_fine", c_i32), ("pairing", c_i32), ("precision", c_i32), ("keep", c_i32)]


class RenderInputs(C.Structure):    # cnrf_render_inputs
    _fields_ = [("ro", C.c_void_p), ("rd", C.c_void_p), ("z_table", C.c_void_p), ("lower", C.c_void_p), ("upper", C.c_void_p),
                ("t_rand", C.c_void_p), ("u", C.c_void_p), ("freqs_xyz", C.c_void_p), ("L_xyz", c_i32),
                ("freqs_dir", C.c_void_p), ("L_dir", c_i32), ("ids", C.c_void_p), ("num_codes", c_i64),
                ("z_s", C.c_void_p), ("ld_zs", c_i64), ("z_t", C.c_void_p), ("ld_zt", c_i64)]


class LossTarget(C.Structure):      # cnrf_loss_target
    _fields_ = [("rows", C.c_void_p), ("ld_rows", c_i64), ("views", C.c_void_p), ("pixel", C.c_void_p), ("channels", c_i32)]


_SIGS = {
    "cnrf_render_loss_fwd": (C.c_int, [C.POINTER(RenderShape), C.POINTER(PtrTable), C.POINTER(PtrTable), C.POINTER(RenderInputs),
                                       C.POINTER(LossTarget), c_f32p, c_f32p, c_f32p, C.c_void_p, c_sz, C.c_void_p, c_sz, C.c_void_p]),
    "cnrf_render_loss_bwd": (C.c_int, [C.POINTER(RenderShape), C.POINTER(PtrTable), C.POINTER(PtrTable), C.POINTER(RenderInputs),
                                       C.POINTER(LossTarget), c_f32p, c_f32p, c_f32p, c_f32p, c_f32p, C.c_void_p, c_sz,
                                       C.POINTER(PtrTable), C.POINTER(PtrTable), c_f32p, c_f32p, C.c_void_p, c_sz, C.c_void_p]),
    "cnrf_render_saved_bytes": (c_sz, [C.POINTER(RenderShape)]),
    "cnrf_render_workspace_bytes": (c_sz, [C.POINTER(RenderShape), c_i32]),
    "cnrf_render_fwd": (C.c_int, [C.POINTER(RenderShape), C.POINTER(PtrTable), C.POINTER(PtrTable), C.POINTER(RenderInputs),
                                  c_f32p, c_f32p, c_f32p, C.c_void_p, c_sz, C.c_void_p, c_sz, C.c_void_p]),
    "cnrf_render_bwd": (C.c_int, [C.POINTER(RenderShape), C.POINTER(PtrTable), C.POINTER(PtrTable), C.POINTER(RenderInputs),
                                  c_f32p, c_f32p, C.c_void_p, c_sz, C.POINTER(PtrTable), C.POINTER(PtrTable), c_f32p, c_f32p,
                                  C.c_void_p, c_sz, C.c_void_p]),
    "cnrf_abi_version": (C.c_int, []),
    "cnrf_error_string": (C.c_char_p, [C.c_int]),
    "cnrf_device_check": (C.c_int, []),
    "cnrf_launch_count": (C.c_int64, []),
    "cnrf_debug_tc_error": (C.c_int, []),
    "cnrf_debug_umma": (C.c_int, [c_i32, c_f32p, c_f32p, c_f32p, c_i32, c_i32, c_f32p, C.c_void_p]),
    "cnrf_debug_bulk_ld": (C.c_int, [c_i32, c_i32, c_i32, c_i32, C.c_void_p, c_sz, c_i32, C.c_void_p, C.c_void_p]),
    "cnrf_debug_bulk_ld2": (C.c_int, [c_i32, c_i32, c_i32, c_i32, c_i32, c_i32, C.c_void_p, c_sz, C.c_void_p, C.c_void_p]),
    "cnrf_u8_to_unit": (C.c_int, [C.c_void_p, c_i64, C.c_void_p, C.c_void_p]),
    "cnrf_gather_pixels_u8": (C.c_int, [C.c_void_p, c_i64, c_i64, c_i32, C.c_void_p, C.c_void_p, c_i64, c_i64, c_i32, C.c_void_p,
                                        C.c_void_p, C.c_void_p]),
    "cnrf_debug_umma_rate": (C.c_int, [c_i32, c_i32, c_i32, c_i32, c_i32, C.c_void_p, c_sz, C.c_void_p, C.c_void_p]),
    "cnrf_debug_tmem_ld": (C.c_int, [c_i32, c_i32, c_i32, c_i32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "cnrf_debug_fwd_trace": (C.c_int, [C.c_void_p]),
    "cnrf_debug_flex_trace": (C.c_int, [C.c_void_p]),
    "cnrf_debug_stream_trace": (C.c_int, [C.c_void_p]),
    "cnrf_debug_fwd_times": (C.c_int, [C.c_void_p, C.c_int]),
    "cnrf_debug_fwd_kernel_ms": (C.c_double, [c_i64, C.c_void_p, C.c_void_p]),
    "cnrf_debug_kernel_ms": (C.c_double, [c_i32, c_i64, C.c_void_p, C.c_void_p]),
    "cnrf_debug_gemm": (C.c_int, [c_i32, c_f32p, c_i64, c_i64, c_f32p, c_i64, c_i64, c_f32p, c_i64, c_i64, c_i32, c_i64, c_i32,
                                  C.c_void_p]),
    "cnrf_ray_bundle_fwd": (C.c_int, [c_f32p, c_i64, c_i32, c_i32, c_f, c_f, c_f, c_f32p, c_i64, c_f32p, c_f32p, C.c_void_p]),
    "cnrf_ray_bundle_bwd": (C.c_int, [c_f32p, c_f32p, c_i64, c_i32, c_i32, c_f, c_f, c_f, c_f32p, c_i64, c_f32p, C.c_void_p]),
    "cnrf_sample_uniform_fwd": (C.c_int, [c_f32p, c_f32p, c_i64, c_i32, c_f32p, c_f32p, c_f32p, c_f32p, c_f32p, c_f32p, C.c_void_p]),
    "cnrf_points_bwd": (C.c_int, [c_f32p, c_f32p, c_i64, c_i32, c_f32p, c_f32p, C.c_void_p]),
    "cnrf_sample_pdf_fwd": (C.c_int, [c_f32p, c_f32p, c_f32p, c_i64, c_f32p, c_i64, c_i32, c_i32, c_f32p, c_f32p, c_f32p, c_f32p, c_f32p, C.c_void_p]),
    "cnrf_invert_cdf": (C.c_int, [c_f32p, c_f32p, c_f32p, c_i64, c_i32, c_i32, c_f32p, c_f32p, C.c_void_p]),
    "cnrf_pos_embed_fwd": (C.c_int, [c_f32p, c_i64, c_i32, c_f32p, c_i32, c_i32, c_f32p, C.c_void_p]),
    "cnrf_pos_embed_bwd": (C.c_int, [c_f32p, c_f32p, c_i64, c_i32, c_f32p, c_i32, c_i32, c_f32p, C.c_void_p]),
    "cnrf_composite_fwd": (C.c_int, [c_f32p, c_f32p, c_f32p, c_i64, c_i32, c_f32p, c_f32p, c_f32p, c_f32p, c_f32p, C.c_void_p]),
    "cnrf_composite_bwd": (C.c_int, [c_f32p, c_f32p, c_f32p, c_i64, c_i32, c_f32p, c_f32p, c_f32p, c_f32p, c_f32p, c_f32p, c_f32p, c_f32p, C.c_void_p]),
    "cnrf_codenerf_saved_bytes": (c_sz, [C.POINTER(FieldShape)]),
    "cnrf_codenerf_workspace_bytes": (c_sz, [C.POINTER(FieldShape), c_i32]),
    "cnrf_field_workspace_bytes": (c_sz, [C.POINTER(FieldShape), c_i32]),
    "cnrf_codenerf_fwd": (C.c_int, [C.POINTER(FieldShape), C.POINTER(PtrTable), c_f32p, c_i64, c_f32p, c_i64, c_f32p, c_i64,
                                    c_f32p, c_i64, c_f32p, C.c_void_p, c_sz, C.c_void_p, c_sz, C.c_void_p]),
    "cnrf_codenerf_bwd": (C.c_int, [C.POINTER(FieldShape), C.POINTER(PtrTable), c_f32p, c_i64, c_f32p, c_i64, c_f32p, c_i64,
                                    c_f32p, c_i64, c_f32p, C.c_void_p, c_sz, C.POINTER(PtrTable), c_f32p, c_f32p, c_f32p,
                                    c_f32p, C.c_void_p, c_sz, C.c_void_p]),
    "cnrf_field_fwd": (C.c_int, [C.POINTER(FieldShape), C.POINTER(PtrTable), c_f32p, c_f32p, c_f32p, c_i32, c_f32p, c_i32,
                                 c_f32p, c_i64, c_f32p, c_i64, c_f32p, C.c_void_p, c_sz, C.c_void_p, c_sz, C.c_void_p]),
    "cnrf_field_fwd_noalloc": (C.c_int, [C.POINTER(FieldShape), C.POINTER(PtrTable), c_f32p, c_f32p, c_f32p, c_i32, c_f32p, c_i32,
                                         c_f32p, c_i64, c_f32p, c_i64, c_f32p, C.c_void_p]),
    "cnrf_field_bwd": (C.c_int, [C.POINTER(FieldShape), C.POINTER(PtrTable), c_f32p, c_f32p, c_f32p, c_i32, c_f32p, c_i32,
                                 c_f32p, c_i64, c_f32p, c_i64, c_f32p, C.c_void_p, c_sz, C.POINTER(PtrTable), c_f32p,
                                 c_f32p, c_f32p, c_f32p, C.c_void_p, c_sz, C.c_void_p]),
    "cnrf_flexnerf_saved_bytes": (c_sz, [C.POINTER(FlexShape)]),
    "cnrf_flexnerf_workspace_bytes": (c_sz, [C.POINTER(FlexShape), c_i32]),
    "cnrf_flexnerf_fwd": (C.c_int, [C.POINTER(FlexShape), C.POINTER(FlexPtrs), c_f32p, c_i64, c_f32p, c_i64, c_f32p, C.c_void_p,
                                    c_sz, C.c_void_p, c_sz, C.c_void_p]),
    "cnrf_flexnerf_fused_inference": (c_i32, [C.POINTER(FlexShape)]),
    "cnrf_flexnerf_fwd_pts": (C.c_int, [C.POINTER(FlexShape), C.POINTER(FlexPtrs), c_f32p, c_i64, c_f32p, c_i32, c_i32, c_f32p,
                                        c_i64, c_f32p, C.c_void_p, c_sz, C.c_void_p]),
    "cnrf_flexnerf_bwd": (C.c_int, [C.POINTER(FlexShape), C.POINTER(FlexPtrs), c_f32p, c_i64, c_f32p, c_i64, c_f32p, C.c_void_p,
                                    c_sz, C.POINTER(FlexPtrs), c_f32p, c_f32p, C.c_void_p, c_sz, C.c_void_p]),
    "cnrf_adamw_step": (C.c_int, [C.c_void_p, c_i32, c_f, c_f, c_f, C.c_void_p]),
    "cnrf_adamw_step_dev": (C.c_int, [C.c_void_p, c_i32, c_f, c_f, c_f, C.c_void_p, C.c_void_p]),
    "cnrf_mse_pair_fwd": (C.c_int, [c_f32p, c_i64, c_f32p, c_i64, c_f32p, c_i64, c_i64, c_f32p, C.c_void_p]),
    "cnrf_mse_pair_bwd": (C.c_int, [c_f32p, c_i64, c_f32p, c_i64, c_f32p, c_i64, c_i64, c_f32p, c_f32p, c_f32p, c_f32p, c_f32p,
                                    C.c_void_p]),
    "cnrf_field_fwd_ids": (C.c_int, [C.POINTER(FieldShape), C.POINTER(PtrTable), c_f32p, c_f32p, c_f32p, c_i32, c_f32p, c_i32,
                                     C.c_void_p, c_i64, c_f32p, c_i64, c_f32p, c_i64, c_f32p, C.c_void_p, c_sz, C.c_void_p,
                                     c_sz, C.c_void_p]),
    "cnrf_field_bwd_ids": (C.c_int, [C.POINTER(FieldShape), C.POINTER(PtrTable), c_f32p, c_f32p, c_f32p, c_i32, c_f32p, c_i32,
                                     C.c_void_p, c_i64, c_f32p, c_i64, c_f32p, c_i64, c_f32p, C.c_void_p, c_sz,
                                     C.POINTER(PtrTable), c_f32p, c_f32p, C.c_void_p, c_sz, C.c_void_p]),
}

EXPORTS = tuple(sorted(_SIGS))
_lib = None


class CnrfError(RuntimeError):
    pass


def load():
    """dlopen libcnrf.so and attach signatures.  Raises if the in-tree library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CnrfError(f"{LIB_PATH} not found: build it with `python code-nerf_b200/build.py` "
                        "(there is no CPU or PyTorch fallback for this path)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in _SIGS.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    if lib.cnrf_abi_version() != 1:
        raise CnrfError("libcnrf.so ABI version mismatch")
    _lib = lib
    if os.environ.get("CODENERF_REPORT_LAUNCHES"):
        # evidence for callers that cannot be edited (the reference's train.py / eval.py run as subprocesses by
        # tests/test_gpu_reference_drivers.py): how many libcnrf kernels this process launched
        import atexit
        atexit.register(lambda: print(f"[libcnrf] pid {os.getpid()}: {lib.cnrf_launch_count()} kernel launches from {LIB_PATH}",
                                      flush=True))
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().cnrf_error_string(int(rc)).decode()
        raise CnrfError(f"{what} failed with code {rc}: {msg}")
