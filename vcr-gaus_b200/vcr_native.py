"""ctypes binding of libvcr_gaus_b200.so (the C ABI declared in include/vcr_gaus_b200.h).

This is the only place the product path touches native code.  There is NO fallback: if the shared
library has not been built (``make -C vcr-gaus_b200`` or ``__graft_entry__.build()``) importing this
module raises, and every wrapper raises ``RuntimeError`` with the library's own message when a call
fails.  PyTorch is used for device memory and streams only.
"""
import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_size_t, c_uint32, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
# VCR_GAUS_LIB lets a developer A/B a differently tuned build of the SAME library; it is not a fallback.
LIB_PATH = os.environ.get("VCR_GAUS_LIB") or os.path.join(_HERE, "libvcr_gaus_b200.so")
if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `make -C {_HERE}` (or __graft_entry__.build()). "
        "There is no CPU or PyTorch fallback for the rasterizer."
    )
lib = ctypes.CDLL(LIB_PATH)

ALLOC_FN = ctypes.CFUNCTYPE(c_void_p, c_void_p, c_size_t)
P_ = c_void_p  # every device pointer travels as void*

_raster_common = [c_int, c_int, c_int]  # P, D, M


def _sig(name, restype, argtypes):
    fn = getattr(lib, name)
    fn.restype = restype
    fn.argtypes = argtypes
    return fn


vcr_abi_version = _sig("vcr_abi_version", c_int, [])
vcr_last_error = _sig("vcr_last_error", c_char_p, [])
vcr_higher_msb = _sig("vcr_higher_msb", c_uint32, [c_uint32])
vcr_geom_bytes = _sig("vcr_geom_bytes", c_size_t, [c_int])
vcr_image_bytes = _sig("vcr_image_bytes", c_size_t, [c_int, c_int])
vcr_binning_bytes = _sig("vcr_binning_bytes", c_size_t, [c_int])
vcr_backward_scratch_bytes = _sig("vcr_backward_scratch_bytes", c_size_t, [c_int])
vcr_knn_scratch_bytes = _sig("vcr_knn_scratch_bytes", c_size_t, [c_int])

vcr_rasterize_forward = _sig(
    "vcr_rasterize_forward", c_int,
    [ALLOC_FN, ALLOC_FN, ALLOC_FN, c_void_p, c_int, c_int, c_int,
     P_, c_int, c_int, P_, P_, P_, P_, P_, P_, P_, c_float, P_, P_, P_, P_, P_, P_,
     c_float, c_float, c_int, P_, P_, P_, c_int, POINTER(c_int), c_void_p])

vcr_rasterize_backward = _sig(
    "vcr_rasterize_backward", c_int,
    [c_int, c_int, c_int, c_int, P_, c_int, c_int, P_, P_, P_, P_, P_, P_, P_, P_, c_float, P_, P_, P_, P_, P_, P_,
     c_float, c_float, P_, P_, P_, P_, P_,
     P_, P_, P_, P_, P_, P_, P_, P_, P_, P_, P_, P_, P_, c_int, c_void_p])

vcr_variant_channels = _sig("vcr_variant_channels", c_int, [c_int, c_int])
vcr_rasterize_forward_variant = _sig("vcr_rasterize_forward_variant", c_int, [c_int, c_int] + list(vcr_rasterize_forward.argtypes))
vcr_rasterize_backward_variant = _sig("vcr_rasterize_backward_variant", c_int, [c_int, c_int] + list(vcr_rasterize_backward.argtypes))

CHUNK_FN = ctypes.CFUNCTYPE(None, c_void_p, c_int, c_int, c_int)   # (user, chunk, first_gaussian, n_gaussians)
vcr_rasterize_backward_chunked = _sig(
    "vcr_rasterize_backward_chunked", c_int,
    [c_int, c_int, c_int, c_int, P_, c_int, c_int, P_, P_, P_, P_, P_, P_, P_, P_, c_float, P_, P_, P_, P_, P_, P_,
     c_float, c_float, P_, P_, P_, P_, P_,
     P_, P_, P_, P_, P_, P_, P_, P_, P_, P_, P_, P_, P_, c_int, c_void_p, c_int, CHUNK_FN, c_void_p, P_])

vcr_multimem_allreduce_sum = _sig("vcr_multimem_allreduce_sum", c_int,
                                  [c_void_p, c_int, POINTER(ctypes.c_longlong), POINTER(ctypes.c_longlong), c_int, c_int, c_int, c_void_p])

vcr_prologue_forward = _sig("vcr_prologue_forward", c_int, [c_int, c_int] + [P_] * 15 + [c_void_p])
vcr_prologue_backward = _sig("vcr_prologue_backward", c_int, [c_int, c_int] + [P_] * 16 + [c_void_p])

vcr_binning_reference_lists = _sig("vcr_binning_reference_lists", c_int, [c_int, c_int, c_int, c_int, P_, P_, P_, c_void_p])
vcr_view_loss_total = _sig("vcr_view_loss_total", c_int, [P_, c_int, c_int, c_float, c_float, c_float, P_, c_void_p])
vcr_unpack_u8 = _sig("vcr_unpack_u8", c_int, [P_, P_, ctypes.c_longlong, c_void_p])
vcr_ssim_forward = _sig("vcr_ssim_forward", c_int, [c_int, c_int, c_int, P_, P_, P_, P_, c_void_p])
vcr_ssim_backward = _sig("vcr_ssim_backward", c_int, [c_int, c_int, c_int, P_, P_, P_, P_, P_, c_void_p])

vcr_adam_step = _sig("vcr_adam_step", c_int,
                     [ctypes.c_longlong, P_, P_, P_, P_, c_int, POINTER(ctypes.c_longlong), POINTER(c_float), c_float, c_float, c_float,
                      c_int, c_float, c_void_p])

vcr_sh_grad_from_colors = _sig("vcr_sh_grad_from_colors", c_int, [c_int, c_int, c_int, c_int, P_, P_, ctypes.c_longlong, P_, c_void_p])
vcr_multimem_exchange = _sig("vcr_multimem_exchange", c_int,
                             [c_void_p, P_, c_int, POINTER(ctypes.c_longlong), POINTER(ctypes.c_longlong), ctypes.c_longlong,
                              ctypes.c_longlong, c_int, c_int, c_int, c_void_p])

vcr_compact_scratch_bytes = _sig("vcr_compact_scratch_bytes", c_size_t, [c_int])
vcr_compact_rows = _sig("vcr_compact_rows", c_int, [c_int, P_, c_int, POINTER(c_void_p), POINTER(c_void_p), POINTER(c_int), P_, P_, c_void_p])

vcr_densify_scratch_bytes = _sig("vcr_densify_scratch_bytes", c_size_t, [c_int])
vcr_densify_decide = _sig("vcr_densify_decide", c_int, [c_int, P_, P_, P_, P_, c_float, c_float, c_float, c_float, P_, P_, P_, c_void_p])
vcr_densify_fill = _sig("vcr_densify_fill", c_int, [c_int, P_, P_, c_int, c_int, c_int, POINTER(c_void_p), POINTER(c_void_p), POINTER(c_int), c_void_p])
vcr_densify_stats_accumulate = _sig("vcr_densify_stats_accumulate", c_int, [c_int, P_, P_, P_, P_, P_, c_void_p])

vcr_mark_visible = _sig("vcr_mark_visible", c_int, [c_int, P_, P_, P_, P_, c_void_p])

vcr_rasterize_count = _sig(
    "vcr_rasterize_count", c_int,
    [ALLOC_FN, ALLOC_FN, ALLOC_FN, c_void_p, c_int, c_int, c_int, P_, c_int, c_int,
     P_, P_, P_, P_, P_, c_float, P_, P_, P_, P_, P_, c_float, c_float, c_int,
     P_, P_, P_, P_, c_int, POINTER(c_int), c_void_p])

vcr_rasterize_visi_acc = _sig(
    "vcr_rasterize_visi_acc", c_int,
    [ALLOC_FN, ALLOC_FN, ALLOC_FN, c_void_p, c_int, c_int, c_int, c_int, c_int,
     P_, P_, P_, P_, P_, c_float, P_, P_, P_, P_, P_, c_float, c_float, c_int,
     P_, P_, c_int, POINTER(c_int), c_void_p])

vcr_dnormal_forward = _sig(
    "vcr_dnormal_forward", c_int,
    [P_, P_, P_, P_, c_int, c_int, c_float, c_float, c_float, c_float, c_float, c_float, P_, P_, P_, c_int, c_void_p])
vcr_dnormal_backward = _sig(
    "vcr_dnormal_backward", c_int,
    [P_, P_, P_, P_, c_int, c_int, c_float, c_float, c_float, c_float, c_float, c_float, P_, P_, P_, P_, c_void_p])
vcr_view_loss_forward = _sig(
    "vcr_view_loss_forward", c_int,
    [P_, P_, P_, P_, c_int, c_int, c_float, c_float, c_float, c_float, c_float, c_float, P_, c_void_p])
vcr_view_loss_backward = _sig(
    "vcr_view_loss_backward", c_int,
    [P_, P_, P_, P_, c_int, c_int, c_float, c_float, c_float, c_float, c_float, c_float, P_, P_, c_float, c_float, c_float, P_,
     c_void_p])
vcr_dist2_knn3 = _sig("vcr_dist2_knn3", c_int, [c_int, P_, P_, P_, c_void_p])
vcr_knn_noalloc_max_points = _sig("vcr_knn_noalloc_max_points", c_int, [])
vcr_dist2_knn3_noalloc = _sig("vcr_dist2_knn3_noalloc", c_int, [c_int, P_, P_, c_void_p])


vcr_pair_counters = _sig("vcr_pair_counters", c_int, [P_])
vcr_debug_block_timeline = _sig("vcr_debug_block_timeline", c_int, [P_])
vcr_debug_radix_fault = _sig("vcr_debug_radix_fault", c_int, [c_int])
vcr_debug_render_version = _sig("vcr_debug_render_version", c_int, [c_int])
vcr_profile_enable = _sig("vcr_profile_enable", c_int, [c_int])
vcr_profile_kernel_count = _sig("vcr_profile_kernel_count", c_int, [])
vcr_profile_kernel_name = _sig("vcr_profile_kernel_name", c_char_p, [c_int])
vcr_profile_read = _sig("vcr_profile_read", c_int, [POINTER(c_float), POINTER(ctypes.c_longlong), POINTER(ctypes.c_longlong), c_int])


def profile_read(reset=True):
    """{kernel name: (total ms over timed launches, timed launches, launches)} since the last reset."""
    n = vcr_profile_kernel_count()
    ms = (c_float * n)()
    timed = (ctypes.c_longlong * n)()
    launches = (ctypes.c_longlong * n)()
    check(vcr_profile_read(ms, timed, launches, int(reset)), "profile_read")
    return {vcr_profile_kernel_name(i).decode(): (float(ms[i]), int(timed[i]), int(launches[i])) for i in range(n)}


class GeomView(ctypes.Structure):
    _fields_ = [(n, c_void_p) for n in ("depths", "clamped", "means2D", "xyz_view", "cov3D", "conic_opacity", "rgb",
                                        "tiles_touched", "point_offsets")]


class BinningView(ctypes.Structure):
    _fields_ = [(n, c_void_p) for n in ("point_list", "point_list_keys", "point_list_unsorted", "point_list_keys_unsorted")]


class ImageView(ctypes.Structure):
    _fields_ = [(n, c_void_p) for n in ("accum_alpha", "n_contrib", "ranges")]


vcr_geom_layout = _sig("vcr_geom_layout", None, [c_void_p, c_int, POINTER(GeomView)])
vcr_binning_layout = _sig("vcr_binning_layout", None, [c_void_p, c_int, POINTER(BinningView)])
vcr_image_layout = _sig("vcr_image_layout", None, [c_void_p, c_int, c_int, POINTER(ImageView)])

# every symbol include/vcr_gaus_b200.h declares (tests check the library exports all of them)
DECLARED_SYMBOLS = [
    "vcr_abi_version", "vcr_last_error", "vcr_higher_msb", "vcr_geom_bytes", "vcr_image_bytes", "vcr_binning_bytes",
    "vcr_rasterize_forward", "vcr_backward_scratch_bytes", "vcr_rasterize_backward", "vcr_mark_visible",
    "vcr_rasterize_count", "vcr_rasterize_visi_acc", "vcr_dnormal_forward", "vcr_dnormal_backward",
    "vcr_knn_scratch_bytes", "vcr_dist2_knn3", "vcr_geom_layout", "vcr_binning_layout", "vcr_image_layout",
    "vcr_knn_noalloc_max_points", "vcr_dist2_knn3_noalloc", "vcr_view_loss_forward", "vcr_view_loss_backward",
    "vcr_profile_enable", "vcr_profile_kernel_count", "vcr_profile_kernel_name", "vcr_profile_read",
    "vcr_rasterize_backward_chunked", "vcr_densify_stats_accumulate", "vcr_multimem_allreduce_sum", "vcr_prologue_forward", "vcr_prologue_backward", "vcr_ssim_forward", "vcr_ssim_backward", "vcr_adam_step", "vcr_view_loss_total", "vcr_unpack_u8", "vcr_binning_reference_lists", "vcr_compact_scratch_bytes", "vcr_compact_rows", "vcr_sh_grad_from_colors", "vcr_multimem_exchange",
    "vcr_variant_channels", "vcr_rasterize_forward_variant", "vcr_rasterize_backward_variant", "vcr_pair_counters", "vcr_debug_block_timeline", "vcr_debug_radix_fault", "vcr_debug_render_version", "vcr_densify_scratch_bytes", "vcr_densify_decide", "vcr_densify_fill",
]


def check(status, what):
    if status != 0:
        raise RuntimeError(f"{what}: {vcr_last_error().decode(errors='replace')}")


def ptr(t):
    """data_ptr of a tensor, or NULL for None / zero-sized tensors (the reference's convention for absent inputs)."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()
