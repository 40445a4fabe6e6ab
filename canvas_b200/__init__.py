"""canvas_b200 -- B200-native per-pixel frame pipeline of fluggo/Canvas.

The product is canvas_b200/libcanvas_b200.so (hand-written sm_100a kernels behind a C-ABI,
include/canvas_b200.h + include/framework.h).  This package is the thin Python host mirror:
`_abi` (ctypes prototypes), `device` (device-resident frames) and `process` (the
fluggo.media.process type surface on top of them).  Nothing here computes pixels on the CPU.
"""
from ._abi import CanvasError, LIB_PATH  # noqa: F401

__all__ = ["CanvasError", "LIB_PATH"]
