"""oracle/ — TEST INFRASTRUCTURE ONLY.

CPU checkers for the scale / pixel-format hot path:

* ``ref_swscale``  — the genuine FFmpeg libswscale binary (the code ffmpeg's ``scale`` filter runs
  for the reference, worker/tasks.py:1565-1580) loaded through ctypes: the parity anchor and the
  "reference" CPU baseline.
* ``port``         — ``swscale_port.c``, a plain-C restatement of SURVEY.md Appendix A, pinned
  against ``ref_swscale`` and tests/golden/ by tests/test_oracle.py.

Only tests/, ``__graft_entry__.smoke()`` and bench.py's cpu_baseline / ``--impl reference`` legs may
import this package.  Nothing under thinvids_b200/ does, and the product has no CPU fallback.
"""
