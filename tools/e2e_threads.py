"""The four host-pointer calls of a page step from 1-4 host threads, with and
without a barrier between steps (pages in flight), wall clock per page.

    python tools/e2e_threads.py [steps] [pageable]
"""
import ctypes
import os
import queue
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libpillowfight_b200 as pf  # noqa: E402
from libpillowfight_b200 import _capi  # noqa: E402
from libpillowfight_b200.pages import A4_300DPI, synthetic_page  # noqa: E402

lib, chk = _capi.lib, _capi.check
w, h = A4_300DPI
K = int(sys.argv[1]) if len(sys.argv) > 1 else 30
PAGEABLE = len(sys.argv) > 2 and sys.argv[2] == "pageable"     # malloc'ed pages instead of pinned ones
page = synthetic_page(w, h, 0)
n = page.nbytes
NG = 3                                      # Gaussian results in flight (free-running Sobel thread)
if PAGEABLE:
    import numpy as np
    heap = np.empty((4 + NG) * n + 4096, dtype=np.uint8)
    pin = (heap.ctypes.data + 4095) // 4096 * 4096
else:
    pin = lib.pf_b200_host_alloc((4 + NG) * n)
ctypes.memmove(pin, page.ctypes.data, n)


def bm(k):
    b = _capi.PfBitmap()
    b.size.x, b.size.y, b.pixels = w, h, pin + k * n
    return b


b_in, b_ace, b_s, b_c = bm(0), bm(1), bm(2), bm(3)
b_g = [bm(4 + k) for k in range(NG)]
xy = pf.ace_samples(100, w, h, 12345)
xyp = xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int))
ace = lambda i: chk(lib.pf_b200_ace(ctypes.byref(b_in), ctypes.byref(b_ace), 100, xyp, 10.0, 1000.0))
gauss = lambda i: chk(lib.pf_b200_gaussian(ctypes.byref(b_in), ctypes.byref(b_g[i % NG]), 2.0, 5))
sobel = lambda i: chk(lib.pf_b200_sobel(ctypes.byref(b_g[i % NG]), ctypes.byref(b_s)))
canny = lambda i: chk(lib.pf_b200_canny(ctypes.byref(b_in), ctypes.byref(b_c)))


def run(groups, barrier, steps, sobel_follows=None, warm=3):
    """groups: one list of calls per thread.  barrier: every thread finishes step i
    before any starts step i+1.  sobel_follows: Sobel runs on its own thread behind
    Gaussian.  The threads warm up first (a host thread's context in the library
    lives as long as the thread), then all start the timed pages together."""
    nthreads = len(groups)
    bar = threading.Barrier(nthreads) if barrier else None
    gate = threading.Barrier(nthreads + 1)
    ready = queue.Queue()                      # Gaussian results ready for Sobel
    free = threading.Semaphore(NG - 1)         # Gaussian buffers Sobel has released
    errs = []

    def work(t, n):
        for i in range(n):
            for call in groups[t]:
                if sobel_follows and call is gauss:
                    free.acquire()
                    call(i)
                    ready.put(i)
                elif sobel_follows and call is sobel:
                    ready.get()
                    call(i)
                    free.release()
                else:
                    call(i)
            if bar:
                bar.wait()

    def loop(t):
        try:
            work(t, warm)
            gate.wait()
            work(t, steps)
            gate.wait()
        except Exception as exc:                   # noqa: BLE001
            errs.append(exc)
            gate.abort()
            if bar:
                bar.abort()

    threads = [threading.Thread(target=loop, args=(t,)) for t in range(nthreads)]
    for th in threads:
        th.start()
    gate.wait()
    t0 = time.perf_counter()
    gate.wait()
    dt = time.perf_counter() - t0
    for th in threads:
        th.join()
    if errs:
        raise errs[0]
    return dt / steps * 1e3


cases = [
    ("1 thread", [[ace, gauss, sobel, canny]], True, None),
    ("2 threads (ace+canny | gauss>sobel), step at a time", [[ace, canny], [gauss, sobel]], True, None),
    ("3 threads (ace | gauss>sobel | canny), step at a time", [[ace], [gauss, sobel], [canny]], True, None),
    ("2 threads, pages in flight", [[ace, canny], [gauss, sobel]], False, None),
    ("3 threads, pages in flight", [[ace], [gauss, sobel], [canny]], False, None),
    ("4 threads (sobel behind gauss), pages in flight", [[ace], [gauss], [sobel], [canny]], False, True),
    ("3 threads (ace+canny | gauss | sobel), pages in flight", [[ace, canny], [gauss], [sobel]], False, True),
]
for name, groups, barrier, follows in cases:
    best = min(run(groups, barrier, K, follows) for _ in range(2))
    print("%-58s %.3f ms per page = %.0f Mpx/s" % (name, best, w * h / best / 1e3), flush=True)
if not PAGEABLE:
    lib.pf_b200_host_free(pin)
