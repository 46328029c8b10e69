"""Host-side mirror of the reference's stereo helper module for the hot path:
stereo_vision/stereo_vision.py:132-184 (compute_disparity) and :187-210 (project_points_3D),
plus parse_config_file (disparity_calculation.py:75-128).  Same names, argument meaning and
error behaviour; the cv2 calls inside are replaced by the CUDA engine.
"""
import configparser
import os

import numpy as np
import torch

from . import engine, ops
from ._lib import error  # noqa: F401  (re-exported)


def parse_config_file(settings_file):
    """ini -> settings dict, defaults and key handling as disparity_calculation.py:75-128.

    One key is added for this engine: `sgbm_mode` ("SGBM", "HH" or "HH4", default "SGBM").  The
    reference's own `mode` key is the phone orientation, not the matcher mode, and is kept
    as such.
    """
    settings = {
        'npz_file': '/media/hao/DATA/Arkit/05122017-105828/src/tmp.npz',
        'mode': 'P',
        'window_size': 3, 'min_disparity': 0, 'num_disparities': 160, 'block_size': 5, 'disp12_max_diff': 1,
        'uniqueness_ratio': 15, 'speckle_window_size': 0, 'speckle_range': 2, 'pre_filter_cap': 63,
        'lmbda': 80000, 'sigma': 1.2, 'sgbm_mode': 'SGBM'}
    if settings_file is None or not os.path.isfile(settings_file):
        return settings
    config = configparser.ConfigParser()
    config.read(settings_file)
    section = config['disparity']
    casts = {'npz_file': str, 'mode': str, 'sgbm_mode': str, 'sigma': float}
    for key in settings:
        if key in section:
            settings[key] = casts.get(key, int)(section[key])
    return settings


_MODES = {'SGBM': engine.STEREO_SGBM_MODE_SGBM, 'HH': engine.STEREO_SGBM_MODE_HH, 'HH4': engine.STEREO_SGBM_MODE_HH4}


def create_left_matcher(disparity_settings, method="SGBM"):
    """The matcher construction of stereo_vision.py:148-168 (P1 = 8*3*ws^2, P2 = 32*3*ws^2)."""
    p1 = 8 * 3 * disparity_settings['window_size'] ** 2
    p2 = 32 * 3 * disparity_settings['window_size'] ** 2
    if method == "SGBM":
        mode = disparity_settings.get('sgbm_mode', 'SGBM')
        if isinstance(mode, str):
            if mode.upper() not in _MODES:
                raise error("sgbm_mode must be SGBM, HH or HH4, got %r" % mode)
            mode = _MODES[mode.upper()]
        return engine.StereoSGBM_create(minDisparity=disparity_settings['min_disparity'],
                                        numDisparities=disparity_settings['num_disparities'],
                                        blockSize=disparity_settings['block_size'],
                                        P1=p1, P2=p2,
                                        disp12MaxDiff=disparity_settings['disp12_max_diff'],
                                        uniquenessRatio=disparity_settings['uniqueness_ratio'],
                                        speckleWindowSize=disparity_settings['speckle_window_size'],
                                        speckleRange=disparity_settings['speckle_range'],
                                        preFilterCap=disparity_settings['pre_filter_cap'],
                                        mode=mode)
    elif method == "BM":
        return engine.StereoBM_create(numDisparities=disparity_settings['num_disparities'],
                                      blockSize=disparity_settings['block_size'])
    raise RuntimeError('Method not supported')


def createRightMatcher(left_matcher):
    """cv2.ximgproc.createRightMatcher for an SGBM matcher (stereo_vision.py:171): the same
    algorithm over the mirrored range, to be called as compute(right, left).
    cv2.ximgproc is absent from this image, so the factory's parameter rule comes from
    SURVEY.md 8(f)-1 and is unpinned; the negative-range engine underneath is parity-tested."""
    if not isinstance(left_matcher, engine.StereoSGBM):
        raise error("createRightMatcher: only StereoSGBM matchers are supported", 2)
    p = {f: getattr(left_matcher, "get" + f[0].upper() + f[1:])() for f in engine.StereoSGBM._FIELDS}
    p['minDisparity'] = -(p['minDisparity'] + p['numDisparities']) + 1
    p['uniquenessRatio'] = 0
    p['disp12MaxDiff'] = 1000000
    p['speckleWindowSize'] = 0
    return engine.StereoSGBM(**p)


def compute_disparity(gray_l, gray_r, disparity_settings, method="SGBM", return_right=False):
    """Disparity of a rectified grayscale pair (stereo_vision.py:132-184).

    Returns (displ, filtered_img) like the reference.  `displ` is the int16 Q4 map of
    left_matcher.compute and is bit-exact with cv2.  The reference's second output is the
    cv2.ximgproc WLS-filtered map (:171-182); that filter is the next row of the scope table
    (no oracle available in this image) and is not built yet, so
    `filtered_img` is None.  With return_right=True the right-view map of :179 is returned
    as a third element.
    """
    left_matcher = create_left_matcher(disparity_settings, method)
    displ = left_matcher.compute(gray_l, gray_r)
    filtered_img = None
    if return_right:
        dispr = createRightMatcher(left_matcher).compute(gray_r, gray_l)
        return displ, filtered_img, dispr
    return displ, filtered_img


def project_points_3D(disparity, Q):
    """Reprojects a disparity map to 3D (stereo_vision.py:187-210).

    The reference divides by 16 whenever `disparity.dtype is not np.float32` -- an identity
    test that is always true, so float32 input is divided too.  Reproduced
    as is; the division is fused into the reprojection kernel (exact: 1/16 is a power of 2).
    """
    if isinstance(disparity, np.ndarray):
        disparity = disparity.astype(np.float32)
    elif isinstance(disparity, torch.Tensor):
        disparity = disparity.to(torch.float32)
    if isinstance(Q, np.ndarray):
        Q = Q.astype(np.float32)
    elif isinstance(Q, torch.Tensor):
        Q = Q.to(torch.float32)
    return ops.reprojectImageTo3D(disparity, Q, disp_scale=1.0 / 16.0)
