"""ctypes binding of libb2stereo.so (include/b2stereo.h).  There is no fallback: if the
library is missing or CUDA is unavailable the calls raise."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B2S_LIB_PATH") or os.path.join(_HERE, "libb2stereo.so")   # override: A/B builds

OK, E_INVALID, E_UNSUPPORTED, E_CUDA, E_NOMEM = 0, 1, 2, 3, 4
DT_U8, DT_I16, DT_I32, DT_F32, DT_U16, DT_F64 = 0, 1, 2, 3, 4, 5


class error(Exception):
    """Counterpart of cv2.error: raised when an argument violates the matcher's contract."""

    def __init__(self, msg, code=E_INVALID):
        super().__init__(msg)
        self.code = code


class SgbmParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "minDisparity", "numDisparities", "blockSize", "P1", "P2", "disp12MaxDiff",
        "preFilterCap", "uniquenessRatio", "speckleWindowSize", "speckleRange", "mode")]


class BmParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "numDisparities", "blockSize", "preFilterCap", "textureThreshold", "uniquenessRatio",
        "speckleWindowSize", "speckleRange", "minDisparity", "disp12MaxDiff")]


class WlsParams(C.Structure):
    _fields_ = [("roi_x", C.c_int32), ("roi_y", C.c_int32), ("roi_w", C.c_int32), ("roi_h", C.c_int32),
                ("min_disp", C.c_int32), ("use_confidence", C.c_int32), ("lrc_thresh", C.c_int32), ("radius", C.c_int32),
                ("roll_off", C.c_float), ("lambda_", C.c_float), ("sigma", C.c_float), ("num_iter", C.c_int32),
                ("attenuation", C.c_float)]


class SgbmGeometry(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "minX1", "maxX1", "width1", "Dp", "invalid", "P1", "P2", "ftzero",
        "uniquenessRatio", "disp12MaxDiff", "ndirs", "R", "G", "swizzled")]


# every symbol include/b2stereo.h declares: (restype, argtypes)
_vp, _i, _sz = C.c_void_p, C.c_int, C.c_size_t
PROTOTYPES = {
    "b2s_last_error": (C.c_char_p, []),
    "b2s_version": (C.c_char_p, []),
    "b2s_launch_count": (C.c_uint64, []),
    "b2s_sgbm_create": (_i, [C.POINTER(SgbmParams), C.POINTER(_vp)]),
    "b2s_sgbm_set_params": (_i, [_vp, C.POINTER(SgbmParams)]),
    "b2s_sgbm_get_geometry": (_i, [_vp, _i, C.POINTER(SgbmGeometry)]),
    "b2s_sgbm_workspace_bytes": (_sz, [_vp, _i, _i, _i]),
    "b2s_sgbm_compute": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _vp]),
    "b2s_sgbm_compute_cn": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp]),
    "b2s_sgbm_keep_volumes": (_i, [_vp, _i]),
    "b2s_sgbm_debug_volumes": (_i, [_vp, _vp, _vp, _sz, _vp]),
    "b2s_sgbm_overflow_flag": (_i, [_vp, C.POINTER(_i)]),
    "b2s_sgbm_enable_timing": (_i, [_vp, _i]),
    "b2s_sgbm_stage_ms": (_i, [_vp, C.POINTER(C.c_float)]),
    "b2s_sgbm_sweep_ms": (_i, [_vp, C.POINTER(C.c_float)]),
    "b2s_bm_create": (_i, [C.POINTER(BmParams), C.POINTER(_vp)]),
    "b2s_bm_compute": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _vp]),
    "b2s_destroy": (None, [_vp]),
    "b2s_wls_create": (_i, [C.POINTER(_vp)]),
    "b2s_wls_destroy": (None, [_vp]),
    "b2s_wls_filter": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, C.POINTER(WlsParams), _vp, _vp, _vp]),
    "b2s_map2dto3d": (_i, [_vp, _i, _vp, _i, _i, _i, _i, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, _vp, _vp, _vp]),
    "b2s_reproject": (_i, [_vp, _i, _i, _i, _i, C.POINTER(C.c_double), C.c_float, _i, _vp, _vp, _vp]),
    "b2s_remap_u8": (_i, [_vp, _i, _i, _i, _vp, _vp, _i, _i, _i, _vp, _vp]),
    "b2s_bgr2gray": (_i, [_vp, _sz, _vp, _vp]),
    "b2s_remap_gray_u8": (_i, [_vp, _i, _i, _vp, _vp, _i, _i, _i, _vp, _vp]),
    "b2s_pipeline": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp,
                          C.POINTER(C.c_double), C.c_float, _i, _vp, _vp, _vp]),
    "b2s_compact_points_scratch_bytes": (_sz, [_i, _i, _i]),
    "b2s_compact_points": (_i, [_vp, _vp, _i, C.c_double, C.c_double, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "b2s_median3_i16": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "b2s_filter_speckles": (_i, [_vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp]),
    "b2s_filter_speckles_noalloc": (_i, [_vp, _i, _i, _i, _i, _i, _i, _vp]),
}

_lib = None


def load():
    """Loads the CUDA library; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "libb2stereo.so is not built (%s). Run `python -m stereo_match_b200.build` "
                "(needs nvcc); there is no CPU fallback." % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc):
    if rc != OK:
        msg = load().b2s_last_error().decode("utf-8", "replace")
        if rc in (E_INVALID, E_UNSUPPORTED):
            raise error(msg, rc)
        raise RuntimeError("b2stereo: " + msg)
