"""ctypes binding of libgaussmart_b200.so (the C ABI declared in include/gaussmart_b200.h).

There is NO fallback: if the shared library is missing or does not export the ABI, importing the
product path raises.  Nothing here imports oracle/.
"""
import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_float, c_int, c_size_t, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
# GAUSSMART_B200_LIB selects a tuning variant built by `python -m gaussmart_b200.build --tag=... -D...`
LIB_PATH = os.environ.get("GAUSSMART_B200_LIB") or os.path.join(_HERE, "lib", "libgaussmart_b200.so")

GSM_OK = 0
GSM_ERR_INVALID_ARGUMENT = 10001
GSM_ERR_BINNING_OVERFLOW = 10002
GSM_ERR_WORKSPACE_TOO_SMALL = 10003
GSM_SURFEL_RECORD_BYTES = 80
GSM_GRAD_RECORD_FLOATS = 20
GSM_DENSIFY_ROLE_COPY, GSM_DENSIFY_ROLE_MOMENT, GSM_DENSIFY_ROLE_XYZ, GSM_DENSIFY_ROLE_SCALING = 0, 1, 2, 3
ABI_VERSION = 9

EXPORTS = [
    "gsm_last_error", "gsm_abi_version", "gsm_geometry_bytes", "gsm_image_bytes", "gsm_binning_bytes", "gsm_binning_persistent_bytes",
    "gsm_binning_sorted_keys_offset", "gsm_image_ranges_offset", "gsm_image_ncontrib_offset",
    "gsm_forward", "gsm_grad_workspace_bytes", "gsm_backward", "gsm_mark_visible",
    "gsm_knn_workspace_bytes", "gsm_knn_dist2",
    "gsm_stage_count", "gsm_stage_name", "gsm_stage_timing_enable", "gsm_stage_timing_read",
    "gsm_sort_pairs32_temp_bytes", "gsm_sort_pairs32", "gsm_sh_grad_expand",
    "gsm_photometric_loss_workspace_bytes", "gsm_photometric_loss", "gsm_adam_step", "gsm_densification_stats",
    "gsm_surface_regularizer_workspace_bytes", "gsm_surface_regularizer",
    "gsm_status_create", "gsm_status_destroy", "gsm_status_query", "gsm_status_wait", "gsm_fp32_peak_probe",
    "gsm_densify_workspace_bytes", "gsm_densify_plan", "gsm_densify_apply",
    "gsm_sh_grad_expand_strided", "gsm_peer_allreduce", "gsm_peer_allreduce_multimem", "gsm_peer_staging_floats", "gsm_peer_scatter", "gsm_peer_sum_broadcast",
    "gsm_photometric_loss_static_max_elements", "gsm_photometric_loss_static",
]


class RegularizerArgs(Structure):
    _fields_ = [("allmap", c_void_p), ("H", c_int), ("W", c_int), ("ray_matrix", c_float * 9), ("ray_origin", c_float * 3),
                ("view_rotation", c_float * 9), ("depth_ratio", c_float), ("lambda_normal", c_float), ("lambda_dist", c_float),
                ("grad_scale", c_float), ("loss_out", c_void_p), ("dL_dallmap", c_void_p), ("rend_normal", c_void_p),
                ("surf_normal", c_void_p), ("surf_depth", c_void_p), ("workspace", c_void_p), ("workspace_bytes", c_size_t)]


class DensifyArgs(Structure):
    _fields_ = [("P", c_size_t), ("xyz", c_void_p), ("opacity", c_void_p), ("scaling", c_void_p), ("rotation", c_void_p),
                ("xyz_gradient_accum", c_void_p), ("denom", c_void_p), ("grad_threshold", c_float), ("dense_limit", c_float),
                ("min_opacity", c_float), ("world_limit", c_float), ("split_inv", c_float), ("has_size_limit", c_int),
                ("split_count", c_int), ("totals_host", c_void_p)]


class DensifyTensor(Structure):
    _fields_ = [("inp", c_void_p), ("out", c_void_p), ("words", c_int), ("role", c_int)]


class AdamTensor(Structure):
    _fields_ = [("param", c_void_p), ("grad", c_void_p), ("exp_avg", c_void_p), ("exp_avg_sq", c_void_p),
                ("n", c_size_t), ("lr", ctypes.c_double), ("beta1", ctypes.c_double), ("beta2", ctypes.c_double),
                ("eps", ctypes.c_double), ("step", c_int)]


class ForwardArgs(Structure):
    _fields_ = [
        ("P", c_int), ("D", c_int), ("M", c_int), ("width", c_int), ("height", c_int),
        ("tan_fovx", c_float), ("tan_fovy", c_float), ("scale_modifier", c_float),
        ("prefiltered", c_int), ("debug", c_int),
        ("background", c_void_p), ("means3D", c_void_p), ("shs", c_void_p), ("colors_precomp", c_void_p),
        ("opacities", c_void_p), ("scales", c_void_p), ("rotations", c_void_p),
        ("transMat_precomp", c_void_p), ("viewmatrix", c_void_p), ("projmatrix", c_void_p),
        ("campos", c_void_p), ("out_color", c_void_p), ("out_others", c_void_p), ("radii", c_void_p),
        ("geometry", c_void_p), ("geometry_bytes", c_size_t),
        ("image", c_void_p), ("image_bytes", c_size_t),
        ("binning", c_void_p), ("binning_bytes", c_size_t), ("binning_capacity", c_size_t),
        ("num_rendered_host", POINTER(c_int)), ("work_counters", c_void_p),
        ("shs_rest", c_void_p), ("raw_params", c_int), ("status", c_void_p),
    ]


class BackwardArgs(Structure):
    _fields_ = [
        ("P", c_int), ("D", c_int), ("M", c_int), ("width", c_int), ("height", c_int),
        ("tan_fovx", c_float), ("tan_fovy", c_float), ("scale_modifier", c_float), ("debug", c_int),
        ("background", c_void_p), ("means3D", c_void_p), ("shs", c_void_p), ("colors_precomp", c_void_p),
        ("scales", c_void_p), ("rotations", c_void_p), ("transMat_precomp", c_void_p),
        ("viewmatrix", c_void_p), ("projmatrix", c_void_p), ("campos", c_void_p), ("radii", c_void_p),
        ("geometry", c_void_p), ("image", c_void_p), ("binning", c_void_p), ("binning_capacity", c_size_t),
        ("dL_dout_color", c_void_p), ("dL_dout_others", c_void_p),
        ("grad_workspace", c_void_p), ("grad_workspace_bytes", c_size_t),
        ("dL_dmeans2D", c_void_p), ("dL_dcolors", c_void_p), ("dL_dopacity", c_void_p),
        ("dL_dmeans3D", c_void_p), ("dL_dtransMat", c_void_p), ("dL_dsh", c_void_p),
        ("dL_dscales", c_void_p), ("dL_drotations", c_void_p), ("accumulate", c_int),
        ("dL_drgb", c_void_p),
        ("densify_grad_norm", c_void_p), ("densify_visible", c_void_p), ("max_radii", c_void_p),
        ("shs_rest", c_void_p), ("dL_dsh_rest", c_void_p), ("opacities", c_void_p), ("raw_params", c_int),
        ("dL_drgb_peers", c_void_p * 8), ("n_rgb_peers", c_int),
    ]


_lib = None


def load():
    """Load the shared library once; raise loudly if it is absent or the ABI does not match."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"gaussmart_b200: native library not built ({LIB_PATH}). Run `python -m gaussmart_b200.build` "
            "(or __graft_entry__.build()). There is no CPU or PyTorch fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name in EXPORTS:
        if not hasattr(lib, name):
            raise ImportError(f"gaussmart_b200: {LIB_PATH} does not export {name}")
    lib.gsm_last_error.restype = c_char_p
    lib.gsm_abi_version.restype = c_int
    for f in ("gsm_geometry_bytes", "gsm_grad_workspace_bytes", "gsm_knn_workspace_bytes"):
        getattr(lib, f).restype = c_size_t
        getattr(lib, f).argtypes = [c_int]
    lib.gsm_image_bytes.restype = c_size_t
    lib.gsm_image_bytes.argtypes = [c_int, c_int]
    lib.gsm_image_ranges_offset.restype = c_size_t
    lib.gsm_image_ranges_offset.argtypes = [c_int, c_int]
    lib.gsm_image_ncontrib_offset.restype = c_size_t
    lib.gsm_image_ncontrib_offset.argtypes = [c_int, c_int]
    lib.gsm_binning_bytes.restype = c_size_t
    lib.gsm_binning_bytes.argtypes = [c_size_t, c_int, c_int]
    lib.gsm_binning_persistent_bytes.restype = c_size_t
    lib.gsm_binning_persistent_bytes.argtypes = [c_size_t]
    lib.gsm_binning_sorted_keys_offset.restype = c_size_t
    lib.gsm_binning_sorted_keys_offset.argtypes = [c_size_t]
    lib.gsm_forward.restype = c_int
    lib.gsm_forward.argtypes = [POINTER(ForwardArgs), c_void_p]
    lib.gsm_backward.restype = c_int
    lib.gsm_backward.argtypes = [POINTER(BackwardArgs), c_void_p]
    lib.gsm_mark_visible.restype = c_int
    lib.gsm_mark_visible.argtypes = [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]
    lib.gsm_knn_dist2.restype = c_int
    lib.gsm_knn_dist2.argtypes = [c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]
    lib.gsm_stage_count.restype = c_int
    lib.gsm_stage_name.restype = c_char_p
    lib.gsm_stage_name.argtypes = [c_int]
    lib.gsm_stage_timing_enable.restype = c_int
    lib.gsm_stage_timing_enable.argtypes = [c_int]
    lib.gsm_stage_timing_read.restype = c_int
    lib.gsm_stage_timing_read.argtypes = [POINTER(ctypes.c_double), POINTER(ctypes.c_longlong), c_int]
    lib.gsm_sh_grad_expand.restype = c_int
    lib.gsm_sh_grad_expand.argtypes = [c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]
    lib.gsm_sort_pairs32_temp_bytes.restype = c_size_t
    lib.gsm_sort_pairs32_temp_bytes.argtypes = [c_size_t]
    lib.gsm_sort_pairs32.restype = c_int
    lib.gsm_sort_pairs32.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_void_p, c_size_t,
                                     c_void_p]
    lib.gsm_photometric_loss_workspace_bytes.restype = c_size_t
    lib.gsm_photometric_loss_workspace_bytes.argtypes = [c_int, c_int, c_int]
    lib.gsm_photometric_loss.restype = c_int
    lib.gsm_photometric_loss.argtypes = [c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_float, c_void_p, c_void_p,
                                         c_void_p, c_size_t, c_void_p]
    lib.gsm_photometric_loss_static_max_elements.restype = c_size_t
    lib.gsm_photometric_loss_static_max_elements.argtypes = []
    lib.gsm_photometric_loss_static.restype = c_int
    lib.gsm_photometric_loss_static.argtypes = [c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_float, c_void_p, c_void_p,
                                                c_void_p]
    lib.gsm_surface_regularizer_workspace_bytes.restype = c_size_t
    lib.gsm_surface_regularizer_workspace_bytes.argtypes = [c_int, c_int]
    lib.gsm_surface_regularizer.restype = c_int
    lib.gsm_surface_regularizer.argtypes = [POINTER(RegularizerArgs), c_void_p]
    lib.gsm_adam_step.restype = c_int
    lib.gsm_adam_step.argtypes = [POINTER(AdamTensor), c_int, c_void_p]
    lib.gsm_densification_stats.restype = c_int
    lib.gsm_densification_stats.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]
    lib.gsm_fp32_peak_probe.restype = c_int
    lib.gsm_fp32_peak_probe.argtypes = [c_int, c_void_p, c_size_t, POINTER(ctypes.c_double), c_void_p]
    lib.gsm_sh_grad_expand_strided.restype = c_int
    lib.gsm_sh_grad_expand_strided.argtypes = [c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p, c_int, c_void_p]
    lib.gsm_peer_allreduce.restype = c_int
    lib.gsm_peer_allreduce.argtypes = [POINTER(c_void_p), c_int, c_int, c_size_t, c_size_t, c_void_p]
    lib.gsm_peer_staging_floats.restype = c_size_t
    lib.gsm_peer_staging_floats.argtypes = [c_int, c_size_t, c_size_t]
    lib.gsm_peer_scatter.restype = c_int
    lib.gsm_peer_scatter.argtypes = [c_void_p, POINTER(c_void_p), c_int, c_int, c_size_t, c_size_t, c_void_p]
    lib.gsm_peer_sum_broadcast.restype = c_int
    lib.gsm_peer_sum_broadcast.argtypes = [c_void_p, POINTER(c_void_p), c_int, c_int, c_size_t, c_size_t, c_void_p]
    lib.gsm_peer_allreduce_multimem.restype = c_int
    lib.gsm_peer_allreduce_multimem.argtypes = [c_void_p, c_int, c_int, c_size_t, c_size_t, c_void_p]
    lib.gsm_densify_workspace_bytes.restype = c_size_t
    lib.gsm_densify_workspace_bytes.argtypes = [c_size_t]
    lib.gsm_densify_plan.restype = c_int
    lib.gsm_densify_plan.argtypes = [POINTER(DensifyArgs), c_void_p, c_size_t, c_void_p]
    lib.gsm_densify_apply.restype = c_int
    lib.gsm_densify_apply.argtypes = [POINTER(DensifyArgs), c_void_p, POINTER(DensifyTensor), c_int, c_size_t, c_void_p,
                                      c_void_p, c_void_p, c_void_p]
    lib.gsm_status_create.restype = c_int
    lib.gsm_status_create.argtypes = [POINTER(c_void_p)]
    lib.gsm_status_destroy.restype = c_int
    lib.gsm_status_destroy.argtypes = [c_void_p]
    lib.gsm_status_query.restype = c_int
    lib.gsm_status_query.argtypes = [c_void_p]
    lib.gsm_status_wait.restype = c_int
    lib.gsm_status_wait.argtypes = [c_void_p, POINTER(ctypes.c_longlong), POINTER(c_int), POINTER(c_int)]
    if lib.gsm_abi_version() != ABI_VERSION:
        raise ImportError(f"gaussmart_b200: ABI version mismatch ({lib.gsm_abi_version()} != {ABI_VERSION})")
    _lib = lib
    return lib


def stage_timing_enable(on=True):
    return bool(load().gsm_stage_timing_enable(1 if on else 0))


def stage_timing_read(reset=True):
    """{stage name: (summed device ms, launches)} for every stage timed since the last reset."""
    lib = load()
    n = lib.gsm_stage_count()
    ms = (ctypes.c_double * n)()
    calls = (ctypes.c_longlong * n)()
    check(lib.gsm_stage_timing_read(ms, calls, 1 if reset else 0), "stage_timing_read")
    return {lib.gsm_stage_name(i).decode(): (ms[i], calls[i]) for i in range(n)}


def last_error():
    return load().gsm_last_error().decode("utf-8", "replace")


def check(code, what):
    if code != GSM_OK:
        raise RuntimeError(f"gaussmart_b200 {what} failed (code {code}): {last_error()}")
