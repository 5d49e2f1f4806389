import ctypes, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import libpillowfight_b200 as pf
from libpillowfight_b200 import _capi
from libpillowfight_b200.pages import A4_300DPI, synthetic_page
lib, chk = _capi.lib, _capi.check
w, h = A4_300DPI
page = synthetic_page(w, h, 0)
n = page.nbytes
pin = lib.pf_b200_host_alloc(2 * n)
ctypes.memmove(pin, page.ctypes.data, n)
bi, bo = _capi.PfBitmap(), _capi.PfBitmap()
bi.size.x, bi.size.y, bi.pixels = w, h, pin
bo.size.x, bo.size.y, bo.pixels = w, h, pin + n
for i in range(4):
    if i == 3:
        os.environ["PF_B200_TRACE_BANDS"] = "1"
    chk(lib.pf_b200_gaussian(ctypes.byref(bi), ctypes.byref(bo), 2.0, 5))
