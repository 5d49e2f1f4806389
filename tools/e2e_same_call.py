"""One entry point called from 1-4 host threads at once, each thread on its own pinned
pages: per-call time in all (what the link gives when calls run side by side).

    python tools/e2e_same_call.py [gaussian sobel canny ace]
"""
import ctypes, os, sys, threading, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import libpillowfight_b200 as pf
from libpillowfight_b200 import _capi
from libpillowfight_b200.pages import A4_300DPI, synthetic_page
lib, chk = _capi.lib, _capi.check
w, h = A4_300DPI
page = synthetic_page(w, h, 0)
n = page.nbytes
T = 4
pin = lib.pf_b200_host_alloc(2 * T * n)
for t in range(T):
    ctypes.memmove(pin + 2 * t * n, page.ctypes.data, n)
def bm(k):
    b = _capi.PfBitmap(); b.size.x, b.size.y, b.pixels = w, h, pin + k * n; return b
xy = pf.ace_samples(100, w, h, 12345)
xyp = xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int))
def mk(name, t):
    bi, bo = bm(2 * t), bm(2 * t + 1)
    if name == "ace": return lambda: chk(lib.pf_b200_ace(ctypes.byref(bi), ctypes.byref(bo), 100, xyp, 10.0, 1000.0))
    if name == "gaussian": return lambda: chk(lib.pf_b200_gaussian(ctypes.byref(bi), ctypes.byref(bo), 2.0, 5))
    if name == "sobel": return lambda: chk(lib.pf_b200_sobel(ctypes.byref(bi), ctypes.byref(bo)))
    return lambda: chk(lib.pf_b200_canny(ctypes.byref(bi), ctypes.byref(bo)))
def run(name, nt, K=30):
    fns = [mk(name, t) for t in range(nt)]
    gate = threading.Barrier(nt + 1)

    def loop(f):
        for _ in range(3):          # (also creates the thread's context in the library)
            f()
        gate.wait()
        for _ in range(K):
            f()
        gate.wait()
    ths = [threading.Thread(target=loop, args=(f,)) for f in fns]
    for th in ths:
        th.start()
    gate.wait()
    t0 = time.perf_counter()
    gate.wait()
    dt = (time.perf_counter() - t0) / (K * nt) * 1e3
    for th in ths:
        th.join()
    print("%-9s %d threads: %.3f ms per call in all (%.1f GB/s each way)" % (name, nt, dt, n / dt / 1e6), flush=True)


for name in (sys.argv[1:] or ["gaussian", "sobel", "canny", "ace"]):
    for nt in [1, 2, 3, 4]:
        run(name, nt)
