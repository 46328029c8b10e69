"""GPU parity at the full BASELINE.json shapes, against live cv2 4.13.0
where the CPU side finishes in seconds, plus size-independent properties at the largest shape."""
import numpy as np
import pytest
import torch

import stereo_match_b200 as sm
from stereo_match_b200.synth import synth_pair, KITTI_Q, kitti_rect_maps
from vectors import CFG1, CFG2, CFG3, CFG4

pytestmark = pytest.mark.gpu


def test_cfg1_sgbm_640x480(cv2ref):
    l, r = synth_pair(CFG1["w"], CFG1["h"], 64, seed=CFG1["seed"])
    got = sm.StereoSGBM_create(**CFG1["kw"]).compute(l, r)
    want = cv2ref.sgbm_compute(l, r, **CFG1["kw"])
    assert np.array_equal(got, want)
    assert 0.5 < (want > -16).mean() < 0.99          # the vector exercises rejection paths


def test_cfg2_bm_1280x720(cv2ref):
    l, r = synth_pair(CFG2["w"], CFG2["h"], CFG2["D"], seed=CFG2["seed"])
    got = sm.StereoBM_create(numDisparities=CFG2["D"], blockSize=CFG2["bs"]).compute(l, r)
    assert np.array_equal(got, cv2ref.bm_compute(l, r, CFG2["D"], CFG2["bs"]))


def test_cfg3_hh_kitti_with_reprojection(cv2ref):
    l, r = synth_pair(CFG3["w"], CFG3["h"], 128, seed=CFG3["seed"])
    m = sm.StereoSGBM_create(**CFG3["kw"])
    got = m.compute(l, r)
    want = cv2ref.sgbm_compute(l, r, **CFG3["kw"])
    assert np.array_equal(got, want), "%d pixels differ" % (got != want).sum()
    assert not m.overflowed()
    xyz = sm.project_points_3D(got, KITTI_Q)
    wxyz = cv2ref.reproject(want.astype(np.float32) / 16.0, KITTI_Q.astype(np.float32))
    assert np.array_equal(xyz.view(np.uint32), wxyz.view(np.uint32))
    fin = np.isfinite(wxyz)
    assert np.all(np.abs(xyz[fin] - wxyz[fin]) <= 1e-5 * np.maximum(np.abs(wxyz[fin]), 1e-30))   # north_star tolerance


def test_cfg5_chain_batched(cv2ref):
    """remap x2 -> HH -> reprojection on a batch; every pair equals the cv2 chain."""
    mx, my = kitti_rect_maps()
    pairs = [synth_pair(1242, 375, 128, seed=s) for s in (0, 1, 2)]
    L = torch.from_numpy(np.stack([p[0] for p in pairs])).cuda()
    R = torch.from_numpy(np.stack([p[1] for p in pairs])).cuda()
    mxd, myd = torch.from_numpy(mx).cuda(), torch.from_numpy(my).cuda()
    Lr = sm.remap(L.unsqueeze(-1), mxd, myd, sm.INTER_LINEAR).squeeze(-1)
    Rr = sm.remap(R.unsqueeze(-1), mxd, myd, sm.INTER_LINEAR).squeeze(-1)
    disp = sm.StereoSGBM_create(**CFG3["kw"]).compute(Lr, Rr)
    xyz = sm.reprojectImageTo3D(disp, KITTI_Q.astype(np.float32), disp_scale=1.0 / 16.0)
    for k, (l, r) in enumerate(pairs):
        lr, rr = cv2ref.remap(l, mx, my), cv2ref.remap(r, mx, my)
        want = cv2ref.sgbm_compute(lr, rr, **CFG3["kw"])
        assert np.array_equal(disp[k].cpu().numpy(), want)
        wxyz = cv2ref.reproject(want.astype(np.float32) / 16.0, KITTI_Q.astype(np.float32))
        assert np.array_equal(xyz[k].cpu().numpy().view(np.uint32), wxyz.view(np.uint32))


def test_cfg4_hh_middlebury_fullres(cv2ref):
    """2864x1988, D=256: 2 x 2.9 GB of int16 volumes.  cv2 needs ~11 s for this one."""
    l, r = synth_pair(CFG4["w"], CFG4["h"], 256, seed=CFG4["seed"])
    m = sm.StereoSGBM_create(**CFG4["kw"])
    got = m.compute(l, r)
    # size-independent properties first
    assert got.shape == l.shape and got.dtype == np.int16
    assert (got[:, :255] == -16).all()                         # columns < minD + D - 1 are never valid (median reaches one column out)
    assert got.max() < 256 * 16 and ((got == -16) | (got >= 0)).all()
    assert np.array_equal(m.compute(l, r), got)                # idempotent on a reused handle
    want = cv2ref.sgbm_compute(l, r, **CFG4["kw"])
    assert np.array_equal(got, want), "%d pixels differ" % (got != want).sum()


@pytest.mark.parametrize("mode", [2, 3])
def test_cfg3_shape_other_modes(mode):
    """MODE_SGBM_3WAY (4 stripes of 94 rows + 13 rows of overlap) and MODE_HH4 at the KITTI shape, gray and colour,
    against live cv2 (which runs these two modes in parallel: about a second each)."""
    import cv2
    from vectors import color_pair
    kw = dict(CFG3["kw"], mode=mode)
    l, r = synth_pair(CFG3["w"], CFG3["h"], 128, seed=7)
    m = sm.StereoSGBM_create(**kw)
    got = m.compute(l, r)
    want = cv2.StereoSGBM_create(**kw).compute(l, r)
    assert np.array_equal(got, want), "%d pixels differ" % (got != want).sum()
    lc, rc = color_pair(CFG3["w"], CFG3["h"], 128, seed=21)
    kwc = dict(kw, blockSize=3)                                   # colour costs are 3x: stay inside the int16 domain
    mc = sm.StereoSGBM_create(**kwc)
    gotc = mc.compute(lc, rc)
    assert not mc.overflowed()
    wantc = cv2.StereoSGBM_create(**kwc).compute(lc, rc)
    assert np.array_equal(gotc, wantc), "%d pixels differ (colour)" % (gotc != wantc).sum()
    # a batch of colour pairs equals its singles
    L = torch.from_numpy(np.stack([lc, lc[:, ::-1].copy()])).cuda()
    R = torch.from_numpy(np.stack([rc, rc[:, ::-1].copy()])).cuda()
    out = mc.compute(L, R)
    assert np.array_equal(out[0].cpu().numpy(), wantc)
    assert np.array_equal(out[1].cpu().numpy(), cv2.StereoSGBM_create(**kwc).compute(lc[:, ::-1].copy(), rc[:, ::-1].copy()))
