"""G independent groups of four threads (one per call), pages in flight."""
import ctypes, os, queue, sys, threading, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import libpillowfight_b200 as pf
from libpillowfight_b200 import _capi
from libpillowfight_b200.pages import A4_300DPI, synthetic_page
lib, chk = _capi.lib, _capi.check
w, h = A4_300DPI
page = synthetic_page(w, h, 0)
n = page.nbytes
NG, GMAX = 3, 3
pin = lib.pf_b200_host_alloc(GMAX * (4 + NG) * n)
xy = pf.ace_samples(100, w, h, 12345)
xyp = xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int))

def bm(k):
    b = _capi.PfBitmap(); b.size.x, b.size.y, b.pixels = w, h, pin + k * n; return b

def group(g, steps, threads):
    base = g * (4 + NG)
    ctypes.memmove(pin + base * n, page.ctypes.data, n)
    b_in, b_ace, b_s, b_c = bm(base), bm(base + 1), bm(base + 2), bm(base + 3)
    b_g = [bm(base + 4 + k) for k in range(NG)]
    ready, free = queue.Queue(), threading.Semaphore(NG - 1)
    def ace():
        for i in range(steps): chk(lib.pf_b200_ace(ctypes.byref(b_in), ctypes.byref(b_ace), 100, xyp, 10.0, 1000.0))
    def gauss():
        for i in range(steps):
            free.acquire(); chk(lib.pf_b200_gaussian(ctypes.byref(b_in), ctypes.byref(b_g[i % NG]), 2.0, 5)); ready.put(i)
    def sobel():
        for _ in range(steps):
            i = ready.get(); chk(lib.pf_b200_sobel(ctypes.byref(b_g[i % NG]), ctypes.byref(b_s))); free.release()
    def canny():
        for i in range(steps): chk(lib.pf_b200_canny(ctypes.byref(b_in), ctypes.byref(b_c)))
    threads += [threading.Thread(target=f) for f in (ace, gauss, sobel, canny)]

def run(G, steps):
    threads = []
    for g in range(G): group(g, steps, threads)
    t0 = time.perf_counter()
    for t in threads: t.start()
    for t in threads: t.join()
    return (time.perf_counter() - t0) / (G * steps) * 1e3

for G in (1, 2, 3):
    run(G, 3)
    best = min(run(G, 30) for _ in range(2))
    print("%d groups of 4 threads: %.3f ms per page = %.0f Mpx/s" % (G, best, w * h / best / 1e3), flush=True)
