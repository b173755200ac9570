"""ctypes binding of libboofsurf.so (include/boofsurf.h).  There is no CPU fallback: if the library is missing this module
raises at import of `lib()`, and every entry point fails with the library's own error when no CUDA device is present."""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libboofsurf.so")
CSRC = os.path.join(_HERE, "csrc")

BS_OK, BS_ERR_ARG, BS_ERR_CUDA, BS_ERR_CAPACITY, BS_ERR_STATE = 0, 1, 2, 3, 4
PIXEL_F32, PIXEL_U8, PIXEL_S32 = 0, 1, 2


class FhCfg(C.Structure):
    _fields_ = [("detect_threshold", C.c_float), ("extract_radius", C.c_int), ("max_features_per_scale", C.c_int),
                ("initial_sample_size", C.c_int), ("initial_size", C.c_int), ("scales_per_octave", C.c_int),
                ("octaves", C.c_int), ("scale_step_size", C.c_int)]


class OriCfg(C.Structure):
    _fields_ = [("variant", C.c_int), ("object_radius_to_scale", C.c_double), ("radius", C.c_int),
                ("sample_period", C.c_double), ("sample_width", C.c_int), ("weight_sigma", C.c_double),
                ("window_size", C.c_double)]


class DescCfg(C.Structure):
    _fields_ = [("variant", C.c_int), ("width_large_grid", C.c_int), ("width_sub_region", C.c_int),
                ("width_sample", C.c_int), ("use_haar", C.c_int), ("weight_sigma", C.c_double), ("overlap", C.c_int),
                ("sigma_large_grid", C.c_double), ("sigma_sub_region", C.c_double)]


class DeviceView(C.Structure):
    _fields_ = [("d_xy", C.c_void_p), ("d_scale", C.c_void_p), ("d_angle", C.c_void_p), ("d_desc", C.c_void_p),
                ("d_white", C.c_void_p), ("d_frame_offset", C.c_void_p), ("total", C.c_long), ("batch", C.c_int),
                ("dof", C.c_int)]


class BoofSurfError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


# every symbol include/boofsurf.h declares: name -> (restype, argtypes)
_vp, _i, _l, _d = C.c_void_p, C.c_int, C.c_long, C.c_double
_pi = C.POINTER(C.c_int)
SIGNATURES = {
    "bs_fh_cfg_default": (None, [C.POINTER(FhCfg)]),
    "bs_ori_cfg_default": (None, [C.POINTER(OriCfg), _i]),
    "bs_desc_cfg_default": (None, [C.POINTER(DescCfg), _i]),
    "bs_last_error": (C.c_char_p, []),
    "bs_version": (C.c_char_p, []),
    "bs_device_count": (_i, [_pi]),
    "bs_surf_create": (_i, [_i, C.POINTER(FhCfg), C.POINTER(OriCfg), C.POINTER(DescCfg), _i, C.POINTER(_vp)]),
    "bs_surf_destroy": (_i, [_vp]),
    "bs_surf_set_workspace": (_i, [_vp, _i, _l]),
    "bs_surf_set_ways": (_i, [_vp, _i]),
    "bs_surf_set_walk_scratch": (_i, [_vp, _l]),
    "bs_surf_detect": (_i, [_vp, _vp, _i, _i, _i, _i]),
    "bs_surf_create_planar": (_i, [_i, C.POINTER(FhCfg), C.POINTER(OriCfg), C.POINTER(DescCfg), _i, _i, C.POINTER(_vp)]),
    "bs_surf_detect_planar": (_i, [_vp, C.POINTER(_vp), _i, _i, _i, _i, _i]),
    "bs_surf_detect_planar_device": (_i, [_vp, _vp, _l, _l, _i, _i, _i, _i, _vp]),
    "bs_surf_detect_device": (_i, [_vp, _vp, _l, _i, _i, _i, _i, _vp]),
    "bs_surf_detect_host_batch": (_i, [_vp, _vp, _l, _i, _i, _i, _i, _l, _vp, _vp, _vp, _vp, _vp, _vp, C.POINTER(_l)]),
    "bs_host_register": (_i, [_vp, C.c_size_t]),
    "bs_host_unregister": (_i, [_vp]),
    "bs_surf_count": (_i, [_vp, _i, _pi]),
    "bs_surf_total": (_i, [_vp, C.POINTER(_l)]),
    "bs_surf_read": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "bs_surf_dof": (_i, [_vp]),
    "bs_surf_set_timing": (_i, [_vp, _i]),
    "bs_surf_get_timing": (_i, [_vp, _vp, _vp]),
    "bs_surf_view": (_i, [_vp, C.POINTER(DeviceView)]),
    "bs_integral": (_i, [_i, _i, _vp, _i, _i, _i, _vp]),
    "bs_hessian": (_i, [_i, _i, _vp, _i, _i, _i, _i, _vp]),
    "bs_fh_intensity": (_i, [_i, _i, _vp, _i, _i, C.POINTER(FhCfg), _pi, _vp, C.POINTER(C.c_long), _vp, _l]),
    "bs_fh_detect": (_i, [_i, _i, _vp, _i, _i, C.POINTER(FhCfg), _i, _pi, _vp, _vp, _vp]),
    "bs_orientation": (_i, [_i, _i, _vp, _i, _i, C.POINTER(OriCfg), _vp, _i, _vp]),
    "bs_describe": (_i, [_i, _i, _vp, _i, _i, C.POINTER(DescCfg), _vp, _vp, _i, _vp, _vp]),
    "bs_assoc_create": (_i, [_i, _d, _i, C.POINTER(_vp)]),
    "bs_assoc_destroy": (_i, [_vp]),
    "bs_assoc_set_max_fit_error": (_i, [_vp, _d]),
    "bs_assoc_set_mode": (_i, [_vp, _i]),
    "bs_assoc_filter_scores": (_i, [_vp, _vp, _i, _vp, _i, _i, _vp, _vp, _vp]),
    "bs_assoc_run": (_i, [_vp, _vp, _i, _vp, _i, _i, _vp, _vp]),
    "bs_assoc_run_device": (_i, [_vp, _vp, _i, _vp, _i, _i, _vp, _vp, _vp]),
    "bs_assoc_forward_static": (_i, [_i, _vp, _i, _vp, _i, _i, _d, _vp, _vp, _vp]),
    "bs_assoc_forward_device": (_i, [_vp, _vp, _i, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "bs_assoc_merge_columns_device": (_i, [_vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "bs_nonmax": (_i, [_i, _vp, _i, _i, _i, _i, C.c_float, _i, _i, _vp, _vp]),
    "bs_surf_retain": (_i, [_vp, _i, _vp, _vp]),
    "bs_features_count": (_i, [_vp, _vp, _vp]),
    "bs_features_view": (_i, [_vp, _vp]),
    "bs_features_destroy": (_i, [_vp]),
    "bs_assoc_features_device": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _vp]),
    "bs_assoc_forward_only_device": (_i, [_vp, _vp, _i, _vp, _i, _i, _vp, _vp, _vp]),
    "bs_assoc_merge_packed_device": (_i, [_vp, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "bs_assoc_validate_device": (_i, [_vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "bs_assoc_matches": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp, _pi, _vp, _pi, _vp, _pi]),
    "bs_assoc_frames_device": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp]),
    "bs_assoc_frames_batch_device": (_i, [_vp, _vp, _i, _vp, _vp, _vp, _vp, _l, _vp, _vp]),
    "bs_assoc_surf_run": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _i, _i, _vp, _vp]),
    "bs_assoc_surf_run_device": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _i, _i, _vp, _vp, _vp]),
    "bs_assoc_surf_matches": (_i, [_vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _pi, _vp, _pi, _vp, _pi]),
    "bs_launch_count": (_l, []),
}

_LIB = None


def build(force=False):
    """Compile every CUDA source for sm_100a into boofcv_b200/libboofsurf.so (in-tree, travels with gpurun)."""
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", "Makefile"))]
    srcs.append(os.path.join(_HERE, "..", "include", "boofsurf.h"))
    stale = force or not os.path.exists(SO_PATH) or any(os.path.getmtime(s) > os.path.getmtime(SO_PATH) for s in srcs)
    if stale:
        subprocess.check_call(["make", "-C", CSRC, "-j8", "-s"])
    return SO_PATH


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(SO_PATH):
            raise ImportError("libboofsurf.so is not built (run __graft_entry__.build() or make -C boofcv_b200/csrc); "
                              "there is no CPU fallback")
        l = C.CDLL(SO_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _LIB = l
    return _LIB


def check(code):
    if code != BS_OK:
        msg = lib().bs_last_error().decode("utf-8", "replace")
        if code == BS_ERR_ARG:
            raise ValueError(msg)          # IllegalArgumentException on the Java side
        raise BoofSurfError(code, msg)
    return code
