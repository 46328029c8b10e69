"""Host-side mirror of the reference's stereo helper module for the hot path:
stereo_vision/stereo_vision.py:132-184 (compute_disparity) and :187-210 (project_points_3D),
plus parse_config_file (disparity_calculation.py:75-128).  Same names, argument meaning and
error behaviour; the cv2 calls inside are replaced by the CUDA engine.
"""
import collections
import configparser
import os
import threading

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
    """cv2.ximgproc.createRightMatcher (stereo_vision.py:171): the same algorithm over the mirrored disparity range,
    to be called as compute(right, left).  StereoSGBM and StereoBM left matchers, like the original.

    cv2.ximgproc is absent from this image, so the factory's parameter rule -- minDisparity = -(minD + D) + 1, same
    D / blockSize (SGBM: P1, P2, mode, preFilterCap), uniquenessRatio = 0 (BM: textureThreshold = 0 as well),
    disp12MaxDiff = 1000000, speckleWindowSize = 0 -- is restated from the opencv_contrib sources from memory and is
    UNPINNED; the negative-range engines underneath are bit-exact with plain cv2.StereoSGBM / cv2.StereoBM."""
    if isinstance(left_matcher, engine.StereoBM):
        m = engine.StereoBM(numDisparities=left_matcher.getNumDisparities(), blockSize=left_matcher.getBlockSize())
        m.setMinDisparity(-(left_matcher.getMinDisparity() + left_matcher.getNumDisparities()) + 1)
        m.setTextureThreshold(0)
        m.setUniquenessRatio(0)
        m.setDisp12MaxDiff(1000000)
        m.setSpeckleWindowSize(0)
        return m
    if not isinstance(left_matcher, engine.StereoSGBM):
        raise error("createRightMatcher: only StereoSGBM and StereoBM matchers are supported", 2)
    p = {f: getattr(left_matcher, "get" + f[0].upper() + f[1:])() for f in engine.StereoSGBM._FIELDS}
    p['minDisparity'] = -(p['minDisparity'] + p['numDisparities']) + 1
    p['uniquenessRatio'] = 0
    p['disp12MaxDiff'] = 1000000
    p['speckleWindowSize'] = 0
    p['speckleRange'] = 0
    return engine.StereoSGBM(**p)


def compute_disparity(gray_l, gray_r, disparity_settings, method="SGBM", return_right=False):
    """Disparity of a rectified grayscale pair (stereo_vision.py:132-184): returns (displ, filtered_img) like the
    reference, both int16 Q4 maps.

    The reference's order of calls is kept: the right matcher is derived from the left one (:171), then
    createDisparityWLSFilter(left_matcher) (:172) -- which, like the original, switches the left matcher's
    uniqueness / LR / speckle filters off, so `displ` is the raw map the filter wants -- then lambda and sigma from
    the settings (:174-175), both computes (:178-179) and wls_filter.filter(displ, gray_l, None, dispr) (:182).
    `displ` and `dispr` are bit-exact with cv2 for these parameters; the ximgproc factories and the filter itself are
    UNPINNED (no cv2.ximgproc in this image, see wls.py).  With return_right=True `dispr` is returned as a third element.
    """
    left_matcher, right_matcher, wls_filter = _chain_objects(gray_l, disparity_settings, method)
    # One upload, one download: the three calls run on device tensors, and the right-view matcher runs on a side stream
    # next to the left one (a single pair keeps a dozen SMs busy, the two matchers share nothing but their read-only inputs).
    gl, kind = engine._to_device(gray_l, "gray_l")
    gr, _ = engine._to_device(gray_r, "gray_r")
    with torch.cuda.device(gl.device):
        cur = torch.cuda.current_stream()
        side = _side_stream(gl.device)
        side.wait_stream(cur)                       # the uploads above were queued on `cur`
        displ = left_matcher.compute(gl, gr)
        with torch.cuda.stream(side):
            dispr = right_matcher.compute(gr, gl)
        for t in (gl, gr, dispr):
            t.record_stream(side)
        cur.wait_stream(side)
        dispr.record_stream(cur)
        filtered_img = wls_filter.filter(displ, gl, None, dispr)
    out = (engine._from_device(displ, kind), engine._from_device(filtered_img, kind))
    if return_right:
        out += (engine._from_device(dispr, kind),)
    return out


def _side_stream(device):
    streams = getattr(_tls, "side", None)
    if streams is None:
        streams = _tls.side = {}
    key = device.index if device.index is not None else torch.cuda.current_device()
    if key not in streams:
        streams[key] = torch.cuda.Stream(device=device)
    return streams[key]


# The reference builds two matchers and a filter on EVERY call.  Built from scratch, each of them allocates (and at the
# end of the call frees) its device workspace, which costs more than the kernels for a single pair.  The objects carry no
# state from one pair to the next, so the last few sets are kept per host thread (matcher objects are not re-entrant,
# like cv2's) and per device, keyed by everything that defines them.  compute_disparity_cache_clear() drops them.
_CHAIN_KEYS = ('window_size', 'min_disparity', 'num_disparities', 'block_size', 'disp12_max_diff', 'uniqueness_ratio',
               'speckle_window_size', 'speckle_range', 'pre_filter_cap', 'lmbda', 'sigma', 'sgbm_mode')
_CHAIN_CACHE_MAX = 2
_tls = threading.local()


def _chain_objects(gray_l, disparity_settings, method):
    from . import wls
    dev = gray_l.device.index if isinstance(gray_l, torch.Tensor) and gray_l.is_cuda else torch.cuda.current_device()
    try:
        key = (method, dev, tuple((k, disparity_settings.get(k)) for k in _CHAIN_KEYS))
        hash(key)
    except TypeError:
        key = None
    cache = getattr(_tls, "chain", None)
    if cache is None:
        cache = _tls.chain = collections.OrderedDict()
    if key is not None and key in cache:
        cache.move_to_end(key)
        return cache[key]
    left_matcher = create_left_matcher(disparity_settings, method)
    right_matcher = createRightMatcher(left_matcher)
    wls_filter = wls.createDisparityWLSFilter(left_matcher)
    wls_filter.setLambda(disparity_settings['lmbda'])
    wls_filter.setSigmaColor(disparity_settings['sigma'])
    objs = (left_matcher, right_matcher, wls_filter)
    if key is not None:
        cache[key] = objs
        while len(cache) > _CHAIN_CACHE_MAX:
            cache.popitem(last=False)
    return objs


def compute_disparity_cache_clear():
    """Drops the matchers / filters (and their device workspaces) that compute_disparity keeps for this host thread."""
    cache = getattr(_tls, "chain", None)
    if cache:
        cache.clear()


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
