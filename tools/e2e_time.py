"""Wall-clock of each host-pointer entry point on a pinned A4 page."""
import ctypes
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libpillowfight_b200 as pf  # noqa: E402
from libpillowfight_b200 import _capi  # noqa: E402
from libpillowfight_b200.pages import A4_300DPI, synthetic_page  # noqa: E402

lib, chk = _capi.lib, _capi.check
w, h = A4_300DPI
page = synthetic_page(w, h, 0)
n = page.nbytes
pin = lib.pf_b200_host_alloc(2 * n)
ctypes.memmove(pin, page.ctypes.data, n)
bi, bo = _capi.PfBitmap(), _capi.PfBitmap()
bi.size.x, bi.size.y, bi.pixels = w, h, pin
bo.size.x, bo.size.y, bo.pixels = w, h, pin + n
xy = pf.ace_samples(100, w, h, 12345)
xyp = xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int))
calls = {
    "ace": lambda: chk(lib.pf_b200_ace(ctypes.byref(bi), ctypes.byref(bo), 100, xyp, 10.0, 1000.0)),
    "gaussian": lambda: chk(lib.pf_b200_gaussian(ctypes.byref(bi), ctypes.byref(bo), 2.0, 5)),
    "sobel": lambda: chk(lib.pf_b200_sobel(ctypes.byref(bi), ctypes.byref(bo))),
    "canny": lambda: chk(lib.pf_b200_canny(ctypes.byref(bi), ctypes.byref(bo))),
}
for name, fn in calls.items():
    for _ in range(3):
        fn()
    t0 = time.perf_counter()
    for _ in range(20):
        fn()
    ms = (time.perf_counter() - t0) / 20 * 1e3
    print("%-9s %.3f ms per call  (%.1f GB/s of page traffic)" % (name, ms, 2 * n / ms / 1e6))
lib.pf_b200_host_free(pin)
