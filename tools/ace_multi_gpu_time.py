"""pf_b200_ace_multi_gpu: ACE 4960 x 7016 / 400 samples (BASELINE.json config 4) from ONE process
through the C-ABI, host page in, host page out, on 1 .. all visible GPUs; checked against the
one-GPU call."""
import ctypes
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libpillowfight_b200 as pf  # noqa: E402
from libpillowfight_b200 import _capi  # noqa: E402
from libpillowfight_b200.pages import A4_300DPI, A4_600DPI, synthetic_page  # noqa: E402

lib, chk = _capi.lib, _capi.check
w, h = A4_600DPI
small = synthetic_page(*A4_300DPI, page_index=0)
page = np.ascontiguousarray(np.repeat(np.repeat(small, 2, axis=0), 2, axis=1)[:h, :w])
n = page.nbytes
pin = lib.pf_b200_host_alloc(3 * n)
views = [np.ctypeslib.as_array((ctypes.c_uint8 * n).from_address(pin + k * n)).reshape(h, w, 4) for k in range(3)]
views[0][:] = page
bms = []
for k in range(3):
    b = _capi.PfBitmap()
    b.size.x, b.size.y, b.pixels = w, h, pin + k * n
    bms.append(b)
xy = pf.ace_samples(400, w, h, 12345)
xyp = xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int))
chk(lib.pf_b200_ace(ctypes.byref(bms[0]), ctypes.byref(bms[1]), 400, xyp, 10.0, 1000.0))
ref = views[1].copy()
ngpu = pf.device_count()
for g in [x for x in (1, 2, 4, 8) if x <= ngpu]:
    for _ in range(3):
        chk(lib.pf_b200_ace_multi_gpu(ctypes.byref(bms[0]), ctypes.byref(bms[2]), 400, xyp, 10.0, 1000.0, None, g))
    t0 = time.perf_counter()
    reps = 10
    for _ in range(reps):
        chk(lib.pf_b200_ace_multi_gpu(ctypes.byref(bms[0]), ctypes.byref(bms[2]), 400, xyp, 10.0, 1000.0, None, g))
    ms = (time.perf_counter() - t0) / reps * 1e3
    print("%d GPU(s): %.3f ms per page, host to host (%.0f Mpx/s)  equal to the one-GPU result: %s" % (
        g, ms, w * h / ms / 1e3, bool(np.array_equal(views[2], ref))))
lib.pf_b200_host_free(pin)
