"""Randomised soak of the CPU restatement (oracle/) against live cv2, far beyond the committed vectors: every mode,
tiny and saturating penalties, uniquenessRatio up to 99, fully negative ranges, colour, blockSize 1-11; StereoBM with any
window / range / filter; remap / cvtColor / medianBlur / filterSpeckles / reprojectImageTo3D (generators: tools/soak_cases.py, shared with the engine's soak tools/soak_parity.py).  Not collected by
pytest (a script: it takes minutes); tests/test_oracle_vs_cv2.py holds the cases it found as fixed regressions.

    python tests/soak_oracle_vs_cv2.py [n_cases] [seed0] [workers]
"""
import multiprocessing as mp
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


sys.path.insert(0, os.path.join(ROOT, "tools"))


class _OracleOps:
    """The restatement's stage operators under the engine's (= cv2's) names, for soak_cases.ops_failures."""
    @staticmethod
    def remap(src, mx, my, interpolation):
        from oracle import port
        return port.remap(src, mx, my)

    @staticmethod
    def cvtColor(src, code):
        from oracle import port
        return port.bgr2gray(src)

    @staticmethod
    def medianBlur(d, k):
        from oracle import port
        return port.median3(d)

    @staticmethod
    def filterSpeckles(img, newVal, size, diff):
        from oracle import port
        return port.filter_speckles(img, newVal, size, diff), None

    @staticmethod
    def reprojectImageTo3D(d, Q, handleMissingValues=False):
        from oracle import port
        return port.reproject(d, Q, handleMissingValues)


def one_ops(k):
    import cv2
    from soak_cases import ops_failures
    f = ops_failures(k, _OracleOps, cv2)
    return (k, len(f), f, "ops") if f else None


def one(k):
    """k >= 0: StereoSGBM case k // 4 of tools/soak_cases.gen_wide / gen_big / gen_params / gen_struct (k % 4) ; k < 0: StereoBM case
    (-k-1) // 2 of gen_bm / gen_bm_struct"""
    import cv2
    from oracle import port, cv2ref
    from soak_cases import gen_wide, gen_bm
    if k < 0:
        from soak_cases import gen_bm_struct
        D, bs, setters, l, r = (gen_bm, gen_bm_struct)[(-k - 1) % 2]((-k - 1) // 2)
        want = cv2ref.bm_compute(l, r, D, bs, **setters)
        if "speckleWindowSize" in setters and not (want != (setters["minDisparity"] - 1) * 16).any():
            return None           # cv2 returns uninitialised memory when nothing is valid and a speckle window is set
        got = port.bm_compute(l, r, D, bs, **setters)
        n = int((got != want).sum())
        return (k, n, (D, bs, setters), l.shape) if n else None
    from soak_cases import gen_big, gen_params, gen_struct
    c = (gen_wide, gen_big, gen_params, gen_struct)[k % 4](k // 4)
    if c is None:
        return None
    kw, l, r = c
    try:
        want = cv2.StereoSGBM_create(**kw).compute(l, r)
    except cv2.error:
        try:
            port.sgbm_compute(l, r, **kw)
        except ValueError:
            return None
        return (k, -1, kw, l.shape)            # cv2 raised, the restatement did not
    try:
        got = port.sgbm_compute(l, r, **kw)
    except ValueError as ex:
        return (k, -2, (kw, str(ex)), l.shape)  # the restatement raised, cv2 did not
    n = int((got != want).sum())
    return (k, n, kw, l.shape) if n else None


if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 400
    s0 = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    nw = int(sys.argv[3]) if len(sys.argv) > 3 else max(1, (os.cpu_count() or 2) // 2)
    with mp.get_context("fork").Pool(nw) as pool:
        res = [x for x in pool.imap_unordered(one, list(range(s0, s0 + n)) + [-1 - i for i in range(s0, s0 + n)], chunksize=4) if x]
        res += [x for x in pool.imap_unordered(one_ops, range(s0, s0 + n), chunksize=4) if x]
    for k, npx, kw, shape in sorted(res):
        print("FAIL case %d: %d pixels differ %r %r" % (k, npx, kw, shape))
    print("oracle soak: %d StereoSGBM + %d StereoBM + %d stage-operator cases, %d failures" % (n, n, n, len(res)))
    sys.exit(1 if res else 0)
