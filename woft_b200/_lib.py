"""ctypes binding of libwoft_b200.so (include/woft_b200.h).

There is no fallback: if the library is missing, does not export the ABI version this package was written
for, or a call fails, a RuntimeError is raised.  Importing this module does not need a GPU; calling a
compute entry point does.
"""
import ctypes
from ctypes import POINTER, c_char_p, c_float, c_int, c_size_t, c_void_p
from pathlib import Path

LIB_PATH = Path(__file__).resolve().parent / "libwoft_b200.so"
ABI_VERSION = 1

OK, ERR_INVALID_ARGUMENT, ERR_CUDA, ERR_UNSUPPORTED, ERR_WORKSPACE = range(5)
CORR_FP32_SIMT, CORR_TF32, CORR_TF32X3, CORR_BF16, CORR_FP16, CORR_FP16X3 = range(6)
CORR_FMAPS_NHWC = 0x100
CORR_STORE_F16 = 0x200
CORR_FMAPS_F16 = 0x400
H_NONHOMOGENEOUS, H_DLT, H_IRLS_L1, H_IRLS_HUBER = range(4)

CORR_MODES = {"fp32": CORR_FP32_SIMT, "tf32": CORR_TF32, "tf32x3": CORR_TF32X3, "bf16": CORR_BF16, "fp16": CORR_FP16,
              "fp16x3": CORR_FP16X3, "fp16s": CORR_FP16 | CORR_STORE_F16}

_FLOATP_ARRAY = POINTER(c_void_p)


class ConvTcArgs(ctypes.Structure):
    """woft_conv_tc_args (include/woft_b200.h), field for field."""
    _fields_ = [("in_", c_void_p), ("in_stride", ctypes.c_longlong), ("in_f16", c_int),
                ("batch", c_int), ("height", c_int), ("width", c_int), ("cin", c_int), ("cin_pad", c_int),
                ("weight_packed", c_void_p),
                ("cout", c_int), ("cout_pad", c_int), ("kh", c_int), ("kw", c_int), ("pad_h", c_int), ("pad_w", c_int),
                ("box_w", c_int), ("box_h", c_int), ("epilogue", c_int), ("relu", c_int), ("scale", c_float),
                ("bias", c_void_p), ("out", c_void_p), ("out_stride", ctypes.c_longlong), ("out_f16", c_int),
                ("out2", c_void_p), ("out2_stride", ctypes.c_longlong), ("out2_f16", c_int), ("h", c_void_p), ("h_stride", ctypes.c_longlong),
                ("z", c_void_p), ("z_stride", ctypes.c_longlong), ("w1x1", c_void_p), ("head_out", c_void_p),
                ("head_bias", c_float), ("row_period", c_int), ("row_valid", c_int), ("debug_trace", c_void_p), ("debug_flags", c_int), ("halo", c_int), ("pair", c_int), ("pdl", c_int), ("stride", c_int), ("residual", c_void_p),
                ("residual_stride", ctypes.c_longlong), ("stats", c_void_p), ("stats_ctas", c_int)]


CONV_BIAS_ACT, CONV_GRU_ZR, CONV_GRU_Q, CONV_HEAD_MEAN, CONV_FLOW_TAPS = range(5)

# name -> (restype, argtypes); kept in step with include/woft_b200.h (tests/test_abi.py checks both ways)
SIGNATURES = {
    "woft_abi_version": (c_int, []),
    "woft_last_error": (c_char_p, []),
    "woft_corr_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int, c_int]),
    "woft_corr_pyramid": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, _FLOATP_ARRAY, c_void_p,
                                  c_void_p, c_size_t, c_int, c_void_p]),
    "woft_corr_pyramid_static": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, _FLOATP_ARRAY, c_void_p,
                                         c_int, c_void_p]),
    "woft_corr_check": (c_int, [c_void_p, c_void_p]),
    "woft_corr_check_static": (c_int, [c_void_p]),
    "woft_corr_set_trace_buffer": (None, [c_void_p]),
    "woft_corr_lookup": (c_int, [_FLOATP_ARRAY, c_int, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p,
                                 c_int, c_void_p]),
    "woft_corr_lookup_check": (c_int, [c_void_p]),
    "woft_corr_lookup_gather": (c_int, [_FLOATP_ARRAY, c_int, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p,
                                        c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                        c_int, c_void_p]),
    "woft_alt_corr_forward": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                      c_int, c_void_p, c_void_p]),
    "woft_homography_fit": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_float, c_void_p,
                                    c_void_p]),
    "woft_proj_errors": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "woft_warp_perspective_u8": (c_int, [c_void_p, c_int, c_int, c_int, POINTER(ctypes.c_double), c_void_p, c_void_p,
                                         c_void_p]),
    "woft_gru_gates": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "woft_gru_update": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "woft_gru_gates_nhwc": (c_int, [c_void_p, c_void_p, c_void_p, c_int, ctypes.c_longlong, c_int, c_void_p, c_void_p,
                                    c_int, c_void_p]),
    "woft_gru_gates_nhwc_split": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int,
                                          ctypes.c_longlong, c_int, c_void_p, c_void_p, c_int, c_void_p]),
    "woft_gru_update_nhwc": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, ctypes.c_longlong, c_int, c_void_p,
                                     c_int, c_void_p, c_void_p]),
    "woft_concat_nhwc": (c_int, [_FLOATP_ARRAY, POINTER(c_int), c_int, ctypes.c_longlong, c_void_p, c_int, c_int,
                                 c_void_p, c_int, c_int, c_void_p]),
    "woft_concat_nhwc_strided": (c_int, [_FLOATP_ARRAY, POINTER(c_int), POINTER(c_int), c_int, ctypes.c_longlong,
                                         c_void_p, c_int, c_int, c_void_p, c_int, c_int, c_void_p]),
    "woft_flow_head_scratch_bytes": (c_size_t, [c_int, c_int, c_int]),
    "woft_flow_head_step": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_size_t,
                                    c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "woft_flow_head_gather": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                      c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "woft_flow_conv7x7_relu": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "woft_flow_out3x3": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "woft_instance_norm_nhwc_scratch_bytes": (c_size_t, [c_int, c_int]),
    "woft_instance_norm_relu_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_float, c_int, c_void_p, c_void_p, c_size_t,
                                             c_void_p]),
    "woft_instance_norm_relu_add_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_float, c_void_p, c_void_p, c_void_p,
                                                 c_size_t, c_void_p]),
    "woft_instance_norm_relu_nhwc_f16": (c_int, [c_void_p, c_int, c_int, c_int, c_float, c_int, c_void_p, c_void_p, c_void_p,
                                                 c_size_t, c_void_p]),
    "woft_instance_norm_from_sums_f16": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_float, c_int, c_void_p,
                                                 c_void_p, c_void_p, c_size_t, c_void_p]),
    "woft_add_relu": (c_int, [c_void_p, c_void_p, ctypes.c_longlong, c_void_p, c_void_p]),
    "woft_images_to_input": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "woft_stem_conv7x7": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "woft_stem_conv7x7_f16": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "woft_weight_head_conv1": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, ctypes.c_longlong, c_int, c_int, c_int,
                                       c_void_p, c_void_p]),
    "woft_weight_head_conv1_stacked_f16": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, ctypes.c_longlong, c_int, c_int,
                                                   c_int, c_void_p, c_void_p]),
    "woft_bias_act_nhwc": (c_int, [c_void_p, c_void_p, ctypes.c_longlong, c_int, c_float, c_int, c_void_p]),
    "woft_context_split_nhwc": (c_int, [c_void_p, c_void_p, ctypes.c_longlong, c_int, c_void_p, c_void_p, c_void_p, c_int,
                                        c_void_p]),
    "woft_select_scratch_bytes": (c_size_t, [c_int, c_int]),
    "woft_select_correspondences": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_int, c_int,
                                            c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "woft_select_correspondences_local": (c_int, [c_void_p, c_void_p, c_void_p, POINTER(ctypes.c_double), c_int, c_int,
                                                  c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                                  c_size_t, c_void_p]),
    "woft_homography_fit_n": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_void_p,
                                      c_void_p]),
    "woft_inlier_fraction": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_float, c_void_p, c_void_p,
                                     c_void_p]),
    "woft_convex_upsample": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                     c_void_p]),
    "woft_instance_norm_relu": (c_int, [c_void_p, c_int, c_int, c_float, c_int, c_void_p, c_void_p]),
    "woft_conv_tc": (c_int, [POINTER(ConvTcArgs), c_void_p]),
    "woft_flow_conv7x7_relu_f16": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "woft_cast_nhwc_f16": (c_int, [c_void_p, c_int, c_int, ctypes.c_longlong, c_void_p, c_int, c_int, c_void_p, c_int, c_int,
                                   c_void_p]),
    "woft_conv_tc_check": (c_int, [c_void_p]),
}

_lib = None


def load():
    """Load the shared library once and declare every signature.  Raises RuntimeError if it cannot."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m woft_b200.build` (needs nvcc). "
            "woft_b200 has no CPU or PyTorch fallback for its CUDA kernels.")
    lib = ctypes.CDLL(str(LIB_PATH))
    for name, (restype, argtypes) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as exc:
            raise RuntimeError(f"{LIB_PATH} does not export {name}; rebuild it") from exc
        fn.restype, fn.argtypes = restype, argtypes
    if lib.woft_abi_version() != ABI_VERSION:
        raise RuntimeError(f"{LIB_PATH} has ABI version {lib.woft_abi_version()}, expected {ABI_VERSION}; rebuild it")
    _lib = lib
    return lib


def last_error():
    return load().woft_last_error().decode(errors="replace")


def check(status, what):
    """Turn a non-zero status into the exception the reference seam would raise."""
    if status == OK:
        return
    msg = f"{what}: {last_error()} (status {status})"
    if status == ERR_INVALID_ARGUMENT:
        raise AssertionError(msg)  # the reference's shape checks raise AssertionError (least_squares_H.py:158-163)
    raise RuntimeError(msg)


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("woft_b200 kernels run on CUDA tensors only; got a tensor on " + str(t.device))


def ptr(t):
    return c_void_p(t.data_ptr()) if t is not None else c_void_p(0)


def ptr_array(tensors):
    arr = (c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])
    return ctypes.cast(arr, _FLOATP_ARRAY), arr  # keep `arr` alive for the duration of the call


def current_stream(device):
    import torch
    return c_void_p(torch.cuda.current_stream(device).cuda_stream)
