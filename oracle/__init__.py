"""CPU ORACLE -- test infrastructure, never imported by the product package.

Two checkers live here:

* ``oracle.port``  : ctypes binding of ``stereo_oracle.c`` (plain-C restatement of
  SURVEY.md Appendix A).  Exposes per-stage intermediates (C volume, per-direction
  L_r, S volume, pre-median disparity) that cv2 does not.
* ``oracle.cv2ref``: the real thing -- live ``cv2`` 4.13.0 (the third-party
  dependency that the reference calls at stereo_vision/stereo_vision.py:153-178,209).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` legs may import this package.
"""
