"""The drop-in as a C program sees it: pf_b200_* on ordinary (pageable) input
AND output pages."""
import ctypes
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libpillowfight_b200 as pf  # noqa: E402
from libpillowfight_b200 import _capi  # noqa: E402
from libpillowfight_b200.pages import A4_300DPI, synthetic_page  # noqa: E402

w, h = A4_300DPI
page = synthetic_page(w, h, 0)
out = np.empty_like(page)
lib = _capi.lib


def bm(a):
    b = _capi.PfBitmap()
    b.size.x, b.size.y, b.pixels = w, h, a.ctypes.data
    return b


bi, bo = bm(page), bm(out)
xy = pf.ace_samples(100, w, h, 12345)
xyp = xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int))
calls = {"ace": lambda: lib.pf_b200_ace(ctypes.byref(bi), ctypes.byref(bo), 100, xyp, 10.0, 1000.0),
         "gaussian": lambda: lib.pf_b200_gaussian(ctypes.byref(bi), ctypes.byref(bo), 2.0, 5),
         "sobel": lambda: lib.pf_b200_sobel(ctypes.byref(bi), ctypes.byref(bo)),
         "canny": lambda: lib.pf_b200_canny(ctypes.byref(bi), ctypes.byref(bo))}
for name, fn in calls.items():
    for _ in range(3):
        assert fn() == 0
    t0 = time.perf_counter()
    for _ in range(10):
        fn()
    print("%-9s %.2f ms per call (pageable in, pageable out)" % (name, (time.perf_counter() - t0) * 100))
