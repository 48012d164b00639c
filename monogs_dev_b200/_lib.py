"""ctypes binding of libb2gs.so (include/b2gs.h).

The library is the product: if it is missing this module raises at import —
there is no CPU or PyTorch fallback anywhere in the package.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb2gs.so")

c_float_p = C.c_void_p  # raw device pointers travel as integers


class B2gsScene(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("D", C.c_int32), ("M", C.c_int32), ("W", C.c_int32), ("H", C.c_int32),
        ("tan_fovx", C.c_float), ("tan_fovy", C.c_float), ("scale_modifier", C.c_float),
        ("prefiltered", C.c_int32), ("debug", C.c_int32),
        ("bg", C.c_void_p), ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p),
        ("projmatrix_raw", C.c_void_p), ("campos", C.c_void_p), ("means3D", C.c_void_p),
        ("shs", C.c_void_p), ("colors_precomp", C.c_void_p), ("opacities", C.c_void_p),
        ("scales", C.c_void_p), ("rotations", C.c_void_p), ("cov3D_precomp", C.c_void_p),
        ("raw_params", C.c_int32), ("shs_rest", C.c_void_p),
    ]


class B2gsWorkspace(C.Structure):
    _fields_ = [
        ("geom", C.c_void_p), ("geom_bytes", C.c_size_t),
        ("image", C.c_void_p), ("image_bytes", C.c_size_t),
        ("binning", C.c_void_p), ("binning_bytes", C.c_size_t),
    ]


class B2gsForwardOut(C.Structure):
    _fields_ = [
        ("color", C.c_void_p), ("depth", C.c_void_p), ("opacity", C.c_void_p),
        ("radii", C.c_void_p), ("n_touched", C.c_void_p), ("status", C.c_void_p),
        ("fused_loss", C.c_void_p), ("bwd_scratch", C.c_void_p), ("bwd_scratch_bytes", C.c_size_t),
        ("event_after_binning", C.c_void_p), ("tile_counters_clean", C.c_int32),
    ]


class B2gsBackwardIO(C.Structure):
    _fields_ = [
        ("dL_dcolor", C.c_void_p), ("dL_ddepth", C.c_void_p), ("radii", C.c_void_p),
        ("dL_dmeans3D", C.c_void_p), ("dL_dmeans2D", C.c_void_p), ("dL_dopacity", C.c_void_p),
        ("dL_dcolors", C.c_void_p), ("dL_dsh", C.c_void_p), ("dL_dscales", C.c_void_p),
        ("dL_drotations", C.c_void_p), ("dL_dcov3D", C.c_void_p), ("dL_dtau", C.c_void_p),
        ("scratch", C.c_void_p), ("scratch_bytes", C.c_size_t),
        ("accumulate", C.c_int32), ("dL_dsh_rest", C.c_void_p), ("densify_grad_accum", C.c_void_p),
        ("densify_denom", C.c_void_p), ("max_radii2D", C.c_void_p),
        ("scratch_zeroed", C.c_int32), ("defer_tau", C.c_int32), ("iso_reg_weight", C.c_float),
        ("iso_reg_stride", C.c_int32), ("iso_reg_offset", C.c_int32),
        ("stats_gate", C.c_void_p * 3),
    ]


class B2gsLossArgs(C.Structure):
    _fields_ = [
        ("N", C.c_int32), ("mode", C.c_int32), ("alpha", C.c_float), ("rgb_boundary_threshold", C.c_float),
        ("image", C.c_void_p), ("depth", C.c_void_p), ("opacity", C.c_void_p), ("gt_image", C.c_void_p),
        ("gt_depth", C.c_void_p), ("exposure_a", C.c_void_p), ("exposure_b", C.c_void_p),
        ("grad_mask", C.c_void_p),
    ]


class B2gsFusedLoss(C.Structure):
    _fields_ = [("args", B2gsLossArgs), ("dL_dimage", C.c_void_p), ("dL_ddepth", C.c_void_p),
                ("partials", C.c_void_p)]


class B2gsAdamSegment(C.Structure):
    _fields_ = [
        ("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p),
        ("exp_avg_sq", C.c_void_p), ("n", C.c_int64), ("lr", C.c_float),
        ("lr_dev", C.c_void_p),
    ]


class B2gsViewFinalize(C.Structure):
    _fields_ = [
        ("loss_partials", C.c_void_p), ("loss_rows", C.c_int32), ("N", C.c_int32), ("mode", C.c_int32),
        ("alpha", C.c_float), ("exposure_a", C.c_void_p), ("tau_partials", C.c_void_p), ("tau_rows", C.c_int32),
        ("reg_scale", C.c_float), ("status", C.c_void_p), ("loss", C.c_void_p), ("dL_dtau", C.c_void_p),
        ("dL_dexposure", C.c_void_p),
    ]


class B2gsPoseCam(C.Structure):
    _fields_ = [
        ("cam_state", C.c_void_p), ("projmatrix_raw", C.c_void_p), ("rot_delta", C.c_void_p),
        ("trans_delta", C.c_void_p), ("exposure_a", C.c_void_p), ("exposure_b", C.c_void_p),
        ("dL_dtau", C.c_void_p), ("dL_dexposure", C.c_void_p), ("moments", C.c_void_p),
        ("lr_rot", C.c_float), ("lr_trans", C.c_float), ("lr_exposure_a", C.c_float),
        ("lr_exposure_b", C.c_float), ("optimize_pose", C.c_int32), ("optimize_exposure", C.c_int32),
        ("tau_norm", C.c_void_p),
    ]


class B2gsDensifyArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("n_groups", C.c_int32),
        ("param", C.c_void_p * 8), ("exp_avg", C.c_void_p * 8), ("exp_avg_sq", C.c_void_p * 8),
        ("width", C.c_int32 * 8),
        ("xyz", C.c_int32), ("opacity", C.c_int32), ("scaling", C.c_int32), ("rotation", C.c_int32),
        ("grad_accum", C.c_void_p), ("denom", C.c_void_p),
        ("max_grad", C.c_float), ("min_opacity", C.c_float), ("dense_threshold", C.c_float),
        ("world_size_threshold", C.c_float), ("max_screen_size_on", C.c_int32),
        ("N", C.c_int32), ("split_divisor", C.c_float), ("aux", C.c_void_p * 2),
    ]


class B2gsPeerArena(C.Structure):
    _fields_ = [
        ("world", C.c_int32), ("rank", C.c_int32),
        ("param_base", C.POINTER(C.c_void_p)), ("grad_base", C.POINTER(C.c_void_p)),
        ("signal", C.POINTER(C.c_void_p)), ("fail_flag", C.c_void_p),
        ("param_mc", C.c_void_p), ("grad_mc", C.c_void_p),
    ]


# every symbol include/b2gs.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "b2gs_version": (C.c_int, []),
    "b2gs_last_error": (C.c_char_p, []),
    "b2gs_kernel_launches": (C.c_uint64, []),
    "b2gs_profile_enable": (None, [C.c_int]),
    "b2gs_profile_read": (C.c_int, [C.c_char_p, C.c_int]),
    "b2gs_geom_bytes": (C.c_size_t, [C.c_int32]),
    "b2gs_image_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "b2gs_binning_bytes": (C.c_size_t, [C.c_int64, C.c_int32, C.c_int32]),
    "b2gs_backward_scratch_bytes": (C.c_size_t, [C.c_int32]),
    "b2gs_backward_tau_partials": (C.c_void_p, [C.c_void_p, C.c_int32, C.POINTER(C.c_int32)]),
    "b2gs_forward_plan": (C.c_int, [C.POINTER(B2gsScene), C.POINTER(B2gsWorkspace),
                                    C.POINTER(B2gsForwardOut), C.c_void_p, C.POINTER(C.c_int64)]),
    "b2gs_forward_render": (C.c_int, [C.POINTER(B2gsScene), C.POINTER(B2gsWorkspace),
                                      C.POINTER(B2gsForwardOut), C.c_int64, C.c_void_p]),
    "b2gs_binning_capacity": (C.c_int64, [C.c_size_t]),
    "b2gs_forward": (C.c_int, [C.POINTER(B2gsScene), C.POINTER(B2gsWorkspace),
                               C.POINTER(B2gsForwardOut), C.c_void_p]),
    "b2gs_backward": (C.c_int, [C.POINTER(B2gsScene), C.POINTER(B2gsWorkspace), C.c_int64,
                                C.POINTER(B2gsBackwardIO), C.c_void_p]),
    "b2gs_mark_visible": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p]),
    "b2gs_debug_export_binning": (C.c_int, [C.POINTER(B2gsScene), C.POINTER(B2gsWorkspace), C.c_int64,
                                            C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b2gs_debug_export_geom": (C.c_int, [C.POINTER(B2gsScene), C.POINTER(B2gsWorkspace), C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b2gs_debug_export_image": (C.c_int, [C.POINTER(B2gsScene), C.POINTER(B2gsWorkspace), C.c_void_p,
                                          C.c_void_p, C.c_void_p]),
    "b2gs_debug_expf_pair_check": (C.c_int, [C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]),
    "b2gs_knn_bytes": (C.c_size_t, [C.c_int32]),
    "b2gs_knn3_dist2": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_size_t,
                                  C.c_void_p]),
    "b2gs_knn3_dist2_static_max_points": (C.c_int32, []),
    "b2gs_knn3_dist2_static": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
    "b2gs_loss_scratch_bytes": (C.c_size_t, [C.c_int32]),
    "b2gs_loss_forward": (C.c_int, [C.POINTER(B2gsLossArgs), C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "b2gs_loss_backward": (C.c_int, [C.POINTER(B2gsLossArgs), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "b2gs_scale_regulariser": (C.c_int, [C.c_int32, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_size_t, C.c_void_p]),
    "b2gs_pose_update": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b2gs_adam_step_dev": (C.c_int, [C.POINTER(B2gsAdamSegment), C.c_int32, C.c_void_p, C.c_float,
                                     C.c_float, C.c_float, C.c_void_p]),
    "b2gs_adam_step": (C.c_int, [C.POINTER(B2gsAdamSegment), C.c_int32, C.c_int64, C.c_float,
                                 C.c_float, C.c_float, C.c_void_p]),
    "b2gs_set_update_guard": (None, [C.c_void_p]),
    "b2gs_debug_set_counters": (None, [C.c_void_p]),
    "b2gs_set_pdl": (None, [C.c_int]),
    "b2gs_backproject_rgbd": (C.c_int, [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_double, C.c_double, C.c_double, C.c_double,
                                        C.POINTER(C.c_double), C.c_float, C.c_void_p, C.c_void_p,
                                        C.c_void_p, C.c_void_p]),
    "b2gs_adam_step_peer": (C.c_int, [C.POINTER(B2gsAdamSegment), C.c_int32, C.c_void_p, C.c_float,
                                      C.c_float, C.c_float, C.POINTER(B2gsPeerArena), C.c_void_p]),
    "b2gs_adam_step_fused": (C.c_int, [C.POINTER(B2gsAdamSegment), C.c_int32, C.c_void_p, C.c_float,
                                       C.c_float, C.c_float, C.c_int32, C.c_void_p]),
    "b2gs_adam_step_peer_fused": (C.c_int, [C.POINTER(B2gsAdamSegment), C.c_int32, C.c_void_p, C.c_float,
                                            C.c_float, C.c_float, C.POINTER(B2gsPeerArena), C.c_int32,
                                            C.c_void_p]),
    "b2gs_densify_bytes": (C.c_size_t, [C.c_int32]),
    "b2gs_densify_plan": (C.c_int, [C.POINTER(B2gsDensifyArgs), C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]),
    "b2gs_densify_apply": (C.c_int, [C.POINTER(B2gsDensifyArgs), C.c_void_p, C.POINTER(C.c_int64), C.c_void_p,
                                     C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                     C.POINTER(C.c_void_p), C.c_void_p]),
    "b2gs_peer_barrier": (C.c_int, [C.POINTER(B2gsPeerArena), C.c_void_p]),
    "b2gs_peer_barrier_status": (C.c_int, [C.POINTER(B2gsPeerArena), C.c_void_p, C.c_int32, C.c_void_p]),
    "b2gs_memset_async": (C.c_int, [C.c_void_p, C.c_int, C.c_size_t, C.c_void_p]),
    "b2gs_window_finalize": (C.c_int, [C.POINTER(B2gsViewFinalize), C.c_int32, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_void_p]),
    "b2gs_pose_step": (C.c_int, [C.POINTER(B2gsPoseCam), C.c_int32, C.c_void_p, C.c_void_p, C.c_float,
                                 C.c_float, C.c_float, C.c_void_p, C.c_void_p]),
}


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python monogs_dev_b200/build.py` "
            "(there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the .so lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


def check(rc):
    if rc != 0:
        raise RuntimeError(f"libb2gs error {rc}: {lib.b2gs_last_error().decode()}")


def profile_read():
    """{kernel name: (launch count, total ms)} since the last read (needs b2gs_profile_enable(1))."""
    buf = C.create_string_buffer(1 << 16)
    n = lib.b2gs_profile_read(buf, len(buf))
    out = {}
    for line in buf.raw[:n].decode().splitlines():
        name, cnt, ms = line.rsplit(" ", 2)
        out[name] = (int(cnt), float(ms))
    return out


def ptr(t):
    """device pointer of a tensor, or None (NULL) for absent / empty tensors."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()
