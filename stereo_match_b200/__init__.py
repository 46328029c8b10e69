"""stereo_match_b200 -- B200 (sm_100a) stereo disparity engine.

Drop-in for the path that ocean1100/stereo_match hands to cv2.StereoSGBM / cv2.StereoBM /
cv2.reprojectImageTo3D / cv2.remap.  Hand-written CUDA behind a C ABI (include/b2stereo.h);
no CPU fallback, no Triton, no multi-backend dispatch.
"""
from ._lib import error, LIB_PATH
from .engine import (StereoSGBM, StereoBM, StereoSGBM_create, StereoBM_create,
                     STEREO_SGBM_MODE_SGBM, STEREO_SGBM_MODE_HH, STEREO_SGBM_MODE_SGBM_3WAY,
                     STEREO_SGBM_MODE_HH4, INTER_LINEAR, COLOR_BGR2GRAY, CV_32F)
from .ops import map2Dto3D, reprojectImageTo3D, remap, cvtColor, medianBlur, filterSpeckles, remap_gray, pipeline
from .cloud import point_cloud, write_ply
from .wls import DisparityWLSFilter, createDisparityWLSFilter, createDisparityWLSFilterGeneric
from . import wls as ximgproc_wls
from .stereo_vision import (compute_disparity, compute_disparity_cache_clear, project_points_3D, parse_config_file,
                            create_left_matcher, createRightMatcher)

__all__ = [
    "error", "LIB_PATH", "StereoSGBM", "StereoBM", "StereoSGBM_create", "StereoBM_create",
    "STEREO_SGBM_MODE_SGBM", "STEREO_SGBM_MODE_HH", "STEREO_SGBM_MODE_SGBM_3WAY", "STEREO_SGBM_MODE_HH4",
    "INTER_LINEAR", "COLOR_BGR2GRAY", "CV_32F",
    "map2Dto3D", "reprojectImageTo3D", "remap", "cvtColor", "medianBlur", "filterSpeckles", "remap_gray", "pipeline",
    "point_cloud", "write_ply", "DisparityWLSFilter", "createDisparityWLSFilter", "createDisparityWLSFilterGeneric",
    "ximgproc",
    "compute_disparity", "compute_disparity_cache_clear", "project_points_3D", "parse_config_file", "create_left_matcher", "createRightMatcher",
]


class _Ximgproc:
    """Namespace mirroring the three cv2.ximgproc names the reference calls (stereo_vision/stereo_vision.py:171-172,
    disparity_test.py:180-182), so that `cv2.ximgproc.X` becomes `stereo_match_b200.ximgproc.X`."""
    createRightMatcher = staticmethod(createRightMatcher)
    createDisparityWLSFilter = staticmethod(createDisparityWLSFilter)
    createDisparityWLSFilterGeneric = staticmethod(createDisparityWLSFilterGeneric)


ximgproc = _Ximgproc()
