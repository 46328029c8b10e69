"""b2s_filter_speckles_noalloc (include/b2stereo.h; stereo_match_b200/csrc/speckle.cu): the speckle filter with its
labels / counts scratch in a static device array inside the library, for callers that cannot call cudaMalloc.  It has
no counterpart in the reference; it must give exactly what the existing
entry point b2s_filter_speckles gives with caller-provided scratch -- integer work, tolerance 0.

This file is named so that pytest collects it LAST.  It passes on its own in a fresh
process and after the other tests: the built-in scratch keeps the previous call's contents by design, and the checks
below run two calls in a row with different images so that stale contents would show.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from stereo_match_b200 import _lib

pytestmark = pytest.mark.gpu

MAX_PIXELS = 1 << 20          # B2S_SPECKLE_NOALLOC_MAX_PIXELS


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _speckled(seed, shape):
    """int16 Q4 disparity-like maps: smooth regions, invalid holes, many small components of every size."""
    rng = np.random.default_rng(seed)
    img = (rng.integers(0, 40, shape) * 16).astype(np.int16)
    b, h, w = shape
    yy, xx = np.mgrid[0:h, 0:w]
    img += ((yy // 9 + xx // 13) % 5 * 16).astype(np.int16)
    img[rng.random(shape) < 0.35] = -16
    img[:, h // 3: h // 3 + 7, :] = 320                     # one long component that must survive
    return img


def _reference_entry_point(img, newVal, maxSize, maxDiff):
    """The existing entry point, scratch provided by the caller."""
    d = torch.from_numpy(img).cuda()
    b, h, w = d.shape
    labels = torch.empty(b * h * w, dtype=torch.int32, device="cuda")
    counts = torch.empty(b * h * w, dtype=torch.int32, device="cuda")
    _lib.check(_lib.load().b2s_filter_speckles(_ptr(d), h, w, b, newVal, maxSize, maxDiff, _ptr(labels), _ptr(counts), _stream()))
    torch.cuda.synchronize()
    return d.cpu().numpy()


def _no_alloc_entry_point(d, newVal, maxSize, maxDiff):
    b, h, w = d.shape
    _lib.check(_lib.load().b2s_filter_speckles_noalloc(_ptr(d), h, w, b, newVal, maxSize, maxDiff, _stream()))


def test_no_alloc_speckle_filter_equals_existing_entry_point_twice_in_a_row():
    torch.cuda.set_device(0)
    cases = [                                   # (images, maxSpeckleSize, maxDiff): the second call is smaller, so the
        (_speckled(1, (1, 375, 1242)), 100, 32),            # scratch still holds the first call's labels around it
        (_speckled(2, (2, 200, 333)), 40, 16),
        (_speckled(3, (1, 375, 1242)), 400, 0),
    ]
    want = [_reference_entry_point(img, -16, size, diff) for img, size, diff in cases]
    dev = [torch.from_numpy(img).cuda() for img, _, _ in cases]
    _no_alloc_entry_point(dev[0].clone(), -16, 100, 32)     # first call on the device resolves the symbol address
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for d, (_, size, diff) in zip(dev, cases):              # calls in a row on one stream, nothing in between
        _no_alloc_entry_point(d, -16, size, diff)
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    assert free1 >= free0, "the no-alloc entry point took %d bytes of device memory" % (free0 - free1)
    for d, w, (img, _, _) in zip(dev, want, cases):
        got = d.cpu().numpy()
        assert np.array_equal(got, w), "%d pixels differ from b2s_filter_speckles" % (got != w).sum()
        assert (w != img).any() and (w == img).any()        # the filter did remove speckles and did keep the rest
    # the same image again after the others: stale scratch contents must not matter
    again = torch.from_numpy(cases[0][0]).cuda()
    _no_alloc_entry_point(again, -16, cases[0][1], cases[0][2])
    assert np.array_equal(again.cpu().numpy(), want[0])


def test_no_alloc_speckle_filter_matches_cv2_and_rejects_oversize():
    cv2 = pytest.importorskip("cv2")
    img = _speckled(4, (1, 120, 250))
    w = img[0].copy()
    cv2.filterSpeckles(w, -16, 60, 32)
    d = torch.from_numpy(img).cuda()
    _no_alloc_entry_point(d, -16, 60, 32)
    assert np.array_equal(d[0].cpu().numpy(), w)
    exact = torch.full((1, 1024, 1024), 16, dtype=torch.int16, device="cuda")         # exactly the limit: accepted
    _no_alloc_entry_point(exact, -16, 10, 0)
    assert bool((exact == 16).all())
    big = torch.zeros((1, 1025, 1024), dtype=torch.int16, device="cuda")
    assert big.numel() > MAX_PIXELS
    with pytest.raises(_lib.error, match="exceeds the built-in scratch"):
        _no_alloc_entry_point(big, -16, 10, 0)
    with pytest.raises(_lib.error, match="bad argument"):
        _lib.check(_lib.load().b2s_filter_speckles_noalloc(None, 4, 4, 1, -16, 1, 1, _stream()))
