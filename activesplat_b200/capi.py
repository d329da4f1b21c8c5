"""ctypes binding of libactivesplat_b200.so (include/activesplat_b200.h).

The library is built in-tree by activesplat_b200/csrc/build.py.  There is NO fallback: if the
shared object is missing or cannot be loaded, importing this module raises.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# ASB_LIB_PATH: developer override for A/B measurements of kernel variants (csrc/build.py build_variant)
LIB_PATH = os.environ.get("ASB_LIB_PATH") or os.path.join(_HERE, "libactivesplat_b200.so")

AS_OK = 0
AS_ERR_INVALID_ARG = -1
AS_ERR_NO_COLORS = -2
AS_ERR_ALLOC = -3
AS_ERR_CUDA = -4
AS_ERR_CAPACITY = -5
AS_ERR_STATIC_LIMIT = -6

c_float_p = ctypes.c_void_p  # device pointers travel as integers
AllocFn = ctypes.CFUNCTYPE(ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t)


class AsForwardArgs(ctypes.Structure):
    _fields_ = [
        ("P", ctypes.c_int), ("D", ctypes.c_int), ("M", ctypes.c_int), ("channels", ctypes.c_int),
        ("background", ctypes.c_void_p),
        ("width", ctypes.c_int), ("height", ctypes.c_int),
        ("means3D", ctypes.c_void_p), ("shs", ctypes.c_void_p), ("colors_precomp", ctypes.c_void_p),
        ("opacities", ctypes.c_void_p), ("scales", ctypes.c_void_p),
        ("scale_modifier", ctypes.c_float),
        ("rotations", ctypes.c_void_p), ("cov3D_precomp", ctypes.c_void_p),
        ("viewmatrix", ctypes.c_void_p), ("projmatrix", ctypes.c_void_p), ("cam_pos", ctypes.c_void_p),
        ("tan_fovx", ctypes.c_float), ("tan_fovy", ctypes.c_float),
        ("prefiltered", ctypes.c_int),
        ("out_color", ctypes.c_void_p), ("out_depth", ctypes.c_void_p), ("out_opacity", ctypes.c_void_p),
        ("radii", ctypes.c_void_p),
        ("debug", ctypes.c_int),
    ]


class AsDensifyArgs(ctypes.Structure):
    _fields_ = [
        ("width", ctypes.c_int), ("height", ctypes.c_int),
        ("color6", ctypes.c_void_p), ("gt_rgb", ctypes.c_void_p), ("gt_depth", ctypes.c_void_p),
        ("sil_thres", ctypes.c_float),
        ("fx", ctypes.c_float), ("fy", ctypes.c_float), ("cx", ctypes.c_float), ("cy", ctypes.c_float),
        ("c2w", ctypes.c_void_p),
        ("isotropic", ctypes.c_int), ("time_idx", ctypes.c_float),
        ("P", ctypes.c_int), ("capacity", ctypes.c_int),
        ("means3D", ctypes.c_void_p), ("rgb_colors", ctypes.c_void_p), ("unnorm_rotations", ctypes.c_void_p),
        ("logit_opacities", ctypes.c_void_p), ("log_scales", ctypes.c_void_p), ("timestep", ctypes.c_void_p),
        ("scratch", ctypes.c_void_p), ("result", ctypes.c_void_p),
    ]


class AsBackwardArgs(ctypes.Structure):
    _fields_ = [
        ("P", ctypes.c_int), ("D", ctypes.c_int), ("M", ctypes.c_int), ("R", ctypes.c_int),
        ("channels", ctypes.c_int),
        ("background", ctypes.c_void_p),
        ("width", ctypes.c_int), ("height", ctypes.c_int),
        ("means3D", ctypes.c_void_p), ("shs", ctypes.c_void_p), ("colors_precomp", ctypes.c_void_p),
        ("scales", ctypes.c_void_p),
        ("scale_modifier", ctypes.c_float),
        ("rotations", ctypes.c_void_p), ("cov3D_precomp", ctypes.c_void_p),
        ("viewmatrix", ctypes.c_void_p), ("projmatrix", ctypes.c_void_p), ("cam_pos", ctypes.c_void_p),
        ("tan_fovx", ctypes.c_float), ("tan_fovy", ctypes.c_float),
        ("radii", ctypes.c_void_p),
        ("geom_buffer", ctypes.c_void_p), ("binning_buffer", ctypes.c_void_p), ("image_buffer", ctypes.c_void_p),
        ("dL_dpix", ctypes.c_void_p),
        ("dL_dmean2D", ctypes.c_void_p), ("dL_dconic", ctypes.c_void_p), ("dL_dopacity", ctypes.c_void_p),
        ("dL_dcolor", ctypes.c_void_p), ("dL_dmean3D", ctypes.c_void_p), ("dL_dcov3D", ctypes.c_void_p),
        ("dL_dsh", ctypes.c_void_p), ("dL_dscale", ctypes.c_void_p), ("dL_drot", ctypes.c_void_p),
        ("debug", ctypes.c_int),
    ]


class AsFrameArgs(ctypes.Structure):
    _fields_ = [
        ("P", ctypes.c_int), ("width", ctypes.c_int), ("height", ctypes.c_int), ("isotropic", ctypes.c_int),
        ("means3D", ctypes.c_void_p), ("rgb_colors", ctypes.c_void_p), ("unnorm_rotations", ctypes.c_void_p),
        ("logit_opacities", ctypes.c_void_p), ("log_scales", ctypes.c_void_p),
        ("cam_unnorm_rot", ctypes.c_void_p), ("cam_rot_stride", ctypes.c_int),
        ("cam_trans", ctypes.c_void_p), ("cam_trans_stride", ctypes.c_int),
        ("sil_w2c", ctypes.c_void_p), ("viewmatrix", ctypes.c_void_p), ("projmatrix", ctypes.c_void_p),
        ("background", ctypes.c_void_p),
        ("tan_fovx", ctypes.c_float), ("tan_fovy", ctypes.c_float), ("scale_modifier", ctypes.c_float),
        ("geom_buffer", ctypes.c_void_p), ("image_buffer", ctypes.c_void_p), ("binning_buffer", ctypes.c_void_p),
        ("binning_capacity", ctypes.c_int),
        ("status", ctypes.c_void_p),
        ("out_color", ctypes.c_void_p), ("out_depth", ctypes.c_void_p), ("out_opacity", ctypes.c_void_p),
        ("radii", ctypes.c_void_p),
        ("debug", ctypes.c_int),
        ("max_2D_radius", ctypes.c_void_p), ("seen", ctypes.c_void_p),
        ("binning_mode", ctypes.c_int),
        ("chunk_scratch", ctypes.c_void_p),
    ]


class AsFrameBackwardArgs(ctypes.Structure):
    _fields_ = [
        ("frame", ctypes.POINTER(AsFrameArgs)),
        ("dL_dpix", ctypes.c_void_p), ("grad_scale", ctypes.c_void_p),
        ("gaussians_grad", ctypes.c_int), ("camera_grad", ctypes.c_int),
        ("grad_params", ctypes.c_void_p), ("grad_means2D", ctypes.c_void_p), ("grad_cam", ctypes.c_void_p),
        ("pose_scratch", ctypes.c_void_p),
        ("accumulate", ctypes.c_int),
        ("grad_cam_stride", ctypes.c_int),
    ]


class AsScratchLayout(ctypes.Structure):
    _fields_ = [(n, ctypes.c_size_t) for n in (
        "depths", "means2D", "conic_opacity", "cov3D", "tiles_touched", "point_offsets", "radii",
        "keys_unsorted", "values_unsorted", "keys_sorted", "point_list",
        "final_T", "n_contrib", "ranges")]


# every symbol include/activesplat_b200.h declares: name -> (restype, argtypes)
_V, _I, _Z, _F = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t, ctypes.c_float
SYMBOLS = {
    "as_forward": (_I, [ctypes.POINTER(AsForwardArgs), AllocFn, AllocFn, AllocFn, _V, _V]),
    "as_backward": (_I, [ctypes.POINTER(AsBackwardArgs), _V]),
    "as_mark_visible": (_I, [_I, _V, _V, _V, _V, _V]),
    "as_geometry_bytes": (_Z, [_I]),
    "as_image_bytes": (_Z, [_I, _I]),
    "as_binning_bytes": (_Z, [_I, _I]),
    "as_scratch_layout": (_I, [_I, _I, _I, _I, _I, ctypes.POINTER(AsScratchLayout)]),
    "as_frame_forward": (_I, [ctypes.POINTER(AsFrameArgs), _V]),
    "as_frame_backward": (_I, [ctypes.POINTER(AsFrameBackwardArgs), _V]),
    "as_loss_scratch_floats": (_Z, [_I, _I]),
    "as_pose_scratch_floats": (_Z, [_I]),
    "as_tile_sort_capacity": (_I, []),
    "as_tracking_loss": (_I, [_I, _I, _V, _V, _V, _F, _F, _F, _I, _I, _V, _V, _V, _V, _V]),
    "as_mapping_loss_scratch_floats": (_Z, [_I, _I]),
    "as_mapping_loss": (_I, [_I, _I, _V, _V, _V, _F, _F, _V, _V, _V, _V, _V]),
    "as_adam_step": (_I, [_I, _I, _V, _V, _V, _V, _V, _V, _V, _V, ctypes.POINTER(ctypes.c_float), _F, _F, _F, _I, _V]),
    "as_chunk_scratch_bytes": (_Z, [_I, _I, _I]),
    "as_densify_scratch_words": (ctypes.c_size_t, [_I, _I]),
    "as_densify": (_I, [ctypes.POINTER(AsDensifyArgs), _V]),
    "as_cluster_hull_volumes": (_I, [_I, _I, _I, _V, _V, _V, ctypes.c_double, _I, _V, _V, _V]),
    "as_adam_step_range": (_I, [_I, _I, _V, _V, _V, _V, _V, _V, _V, _V, ctypes.POINTER(ctypes.c_float), _F, _F, _F, _V,
                                ctypes.c_longlong, ctypes.c_longlong, _I, _V]),
    "as_pose_adam_step": (_I, [_V, _V, _I, _F, _V, _V, _I, _F, _V, _V, _F, _F, _F, _V, _V]),
    "as_adam_step_counted": (_I, [_I, _I, _V, _V, _V, _V, _V, _V, _V, _V, ctypes.POINTER(ctypes.c_float), _F, _F, _F, _V,
                                  _V]),
    "as_enable_peer_access": (_I, [_I]),
    "as_ipc_open": (_I, [ctypes.c_char_p, ctypes.POINTER(ctypes.c_void_p)]),
    "as_ipc_close": (_I, [_V]),
    "as_adam_step_peer": (_I, [_I, _I, _I, _I, _I, _I, _V, _V, _V, _V, ctypes.POINTER(ctypes.c_float), _F, _F, _F, _I, _V]),
    "as_dbscan_grid": (_I, [_I, _I, _I, _V, _F, _F, _I, _V, _V, _I, _V, _V, _V, _V]),
    "as_sort_temp_bytes": (_Z, [_I]),
    "as_sort_pairs": (_I, [_V, _V, _V, _V, _I, _I, _V, _Z, _V]),
    "as_scan_temp_bytes": (_Z, [_I]),
    "as_inclusive_scan_u32": (_I, [_V, _V, _I, _V, _V, _Z, _V]),
    "as_static_scan_max_n": (_I, []),
    "as_inclusive_scan_u32_static": (_I, [_V, _V, _I, _V, _V]),
    "as_higher_msb": (ctypes.c_uint32, [ctypes.c_uint32]),
    "as_profile_enable": (None, [_I]),
    "as_profile_read": (_I, [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int), _I]),
    "as_launch_count": (ctypes.c_longlong, []),
    "as_error_string": (ctypes.c_char_p, [_I]),
    "as_last_cuda_error": (_I, []),
    "as_version": (ctypes.c_char_p, []),
}

STAGES = ["preprocess", "scan", "duplicate", "sort", "ranges_gather", "blend_fwd", "blend_bwd", "preprocess_bwd",
          "prologue", "loss", "adam", "other"]

_lib = None


def lib():
    """Load the shared library (once).  Raises if it is absent -- there is no CPU or eager fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python activesplat_b200/csrc/build.py` "
                "(or __graft_entry__.build()). activesplat_b200 has no fallback path.")
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)  # AttributeError if the library does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


class AsError(RuntimeError):
    pass


def check(code):
    if code < 0:
        raise AsError(lib().as_error_string(code).decode())
    return code


def profile_read():
    """{stage: (ms, calls)} since the last read (see as_profile_read)."""
    n = len(STAGES)
    ms = (ctypes.c_double * n)()
    calls = (ctypes.c_int * n)()
    check(lib().as_profile_read(ms, calls, n))
    return {STAGES[i]: (ms[i], calls[i]) for i in range(n)}
