"""Randomised parity sweep on the GPU: random shapes, contents and parameters
through every entry point against the CPU oracle.
Usage: fuzz_parity.py [cases] [seed] [pinned]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import pf_oracle as o  # noqa: E402

rng = None


def content(w, h):
    kind = rng.integers(0, 6)
    if kind == 0:
        img = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
    elif kind == 1:                                   # flat with a few spikes
        img = np.full((h, w, 4), rng.integers(0, 256), dtype=np.uint8)
        for _ in range(rng.integers(0, 6)):
            img[rng.integers(0, h), rng.integers(0, w), :3] = rng.integers(0, 256, size=3)
    elif kind == 2:                                   # blocks (strong and weak edges)
        img = np.zeros((h, w, 4), dtype=np.uint8)
        for _ in range(rng.integers(1, 12)):
            y0, x0 = rng.integers(0, h), rng.integers(0, w)
            img[y0:y0 + rng.integers(1, h + 1), x0:x0 + rng.integers(1, w + 1), :3] = rng.integers(0, 256, size=3)
    elif kind == 3:                                   # smooth ramp plus noise
        yy, xx = np.mgrid[0:h, 0:w]
        base = (xx * rng.uniform(0, 3) + yy * rng.uniform(0, 3)) % 256
        img = np.stack([base, base[::-1], base.T[:h, :w] if h == w else base, base], -1).astype(np.uint8)
        img[..., :3] = np.clip(img[..., :3].astype(int) + rng.integers(-8, 9, size=(h, w, 3)), 0, 255)
    elif kind == 4:                                   # low-contrast texture (many weak pixels)
        img = (rng.integers(0, 2, size=(h, w, 1)) * rng.integers(20, 90)).astype(np.uint8).repeat(4, -1)
    else:                                             # saturated page
        img = np.full((h, w, 4), 255, dtype=np.uint8)
        img[rng.integers(0, h):, rng.integers(0, w):, :3] = rng.integers(0, 256)
    img[..., 3] = rng.integers(0, 256) if rng.random() < 0.3 else 255
    return np.ascontiguousarray(img)


def gen_cases(cases, seed):
    """Yields (case, img, gaussian (sigma, taps), ace (slope, limit, samples, seed) or None);
    the stream depends on `seed` only, so a case can be replayed on the CPU."""
    global rng
    rng = np.random.default_rng(seed)
    for case in range(cases):
        w = int(rng.integers(1, 700)) if rng.random() < 0.5 else 4 * int(rng.integers(1, 180))
        h = int(rng.integers(1, 500))
        img = content(w, h)
        n = int(rng.choice([1, 2, 3, 4, 5, 5, 5, 6, 7, 9, 11, 15]))
        sigma = float(rng.choice([0.0, 0.7, 1.0, 2.0, 3.5]))
        ace = None
        if h >= 10:
            ns = int(rng.choice([8, 20, 64, 130]))
            slope, limit = [(10.0, 1000.0), (5.0, 500.0), (7.5, 1000.0), (20.0, 100.0)][rng.integers(0, 4)]
            ace = (slope, limit, ns, int(rng.integers(0, 1 << 30)))
        yield case, img, (sigma, n), ace


class _Pinned:
    """The same four calls on pinned caller memory (the band-pipelined paths),
    straight through the C-ABI; tall cases only (>= 512 rows stream)."""

    def __init__(self):
        import ctypes
        from libpillowfight_b200 import _capi
        self.ct, self.capi, self.lib = ctypes, _capi, _capi.lib
        self.cap, self.ptr = 0, None

    def call(self, fn, img, *args):
        ct, n = self.ct, img.nbytes
        if 2 * n > self.cap:
            if self.ptr:
                self.lib.pf_b200_host_free(self.ptr)
            self.cap = 2 * n
            self.ptr = self.lib.pf_b200_host_alloc(self.cap)
        src = np.ctypeslib.as_array((ct.c_uint8 * n).from_address(self.ptr)).reshape(img.shape)
        dst = np.ctypeslib.as_array((ct.c_uint8 * n).from_address(self.ptr + n)).reshape(img.shape)
        src[:] = img
        dst[:] = 0x5A
        bi, bo = self.capi.PfBitmap(), self.capi.PfBitmap()
        bi.size.x = bo.size.x = img.shape[1]
        bi.size.y = bo.size.y = img.shape[0]
        bi.pixels, bo.pixels = self.ptr, self.ptr + n
        self.capi.check(fn(ct.byref(bi), ct.byref(bo), *args))
        return dst.copy()


def run(cases, seed, pinned=False):
    """-> list of mismatch descriptions (empty: parity on every case).
    pinned: tall cases go through the pinned-memory entry points instead."""
    import ctypes
    import libpillowfight_b200 as pf
    bad = []
    pin = _Pinned() if pinned else None
    for case, img, (sigma, n), ace in gen_cases(cases, seed):
        if pinned:
            reps = -(-600 // img.shape[0])             # tile the page to >= 600 rows
            img = np.ascontiguousarray(np.tile(img, (reps, 1, 1)))
        h, w = img.shape[:2]
        if pinned:
            lib = pin.lib
            checks = [("gaussian(%g,%d)" % (sigma, n), lambda: pin.call(lib.pf_b200_gaussian, img, sigma, n),
                       lambda: o.gaussian(img, sigma, n), 0),
                      ("sobel", lambda: pin.call(lib.pf_b200_sobel, img), lambda: o.sobel(img), 0),
                      ("canny", lambda: pin.call(lib.pf_b200_canny, img), lambda: o.canny(img), 0)]
        else:
            checks = [("gaussian(%g,%d)" % (sigma, n), lambda: pf.gaussian(img, sigma, n), lambda: o.gaussian(img, sigma, n), 0),
                      ("sobel", lambda: pf.sobel(img), lambda: o.sobel(img), 0),
                      ("canny", lambda: pf.canny(img), lambda: o.canny(img), 0)]
        if ace:
            slope, limit, ns, seed = ace
            xy = o.ace_draw_samples(ns, w, h, seed)
            if pinned:
                xyp = xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int))
                checks.append(("ace(%g,%g,%d)" % (slope, limit, ns),
                               lambda: pin.call(pin.lib.pf_b200_ace, img, ns, xyp, slope, limit),
                               lambda: o.ace_with_samples(img, xy, slope, limit, nb_threads=8), 0))
            else:
                checks.append(("ace(%g,%g,%d)" % (slope, limit, ns), lambda: pf.ace(img, slope, limit, ns, seed=seed),
                               lambda: o.ace_with_samples(img, xy, slope, limit, nb_threads=4), 0))
        for name, fg, fr, tol in checks:
            got, ref = fg(), fr()
            d = np.abs(got.astype(np.int16) - ref.astype(np.int16))
            frac = float((d[..., :3] != 0).mean())
            ok = d.max() <= tol and (tol == 0 or frac <= 1e-3 or (d != 0).sum() <= 3)
            if not ok:
                bad.append("case %d %dx%d %s: max %d, %.4f %% differ" % (case, w, h, name, d.max(), 100 * frac))
    return bad


def main():
    cases = int(sys.argv[1]) if len(sys.argv) > 1 else 100
    bad = run(cases, int(sys.argv[2]) if len(sys.argv) > 2 else 0, pinned="pinned" in sys.argv[3:])
    for line in bad:
        print("MISMATCH", line)
    print("fuzz: %d cases, %d mismatches" % (cases, len(bad)))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
