"""B200 (sm_100a) build of libpillowfight's hot path, with the call surface of
the reference's `pillowfight` Python module for that path:

    ace(img_in, slope=10, limit=1000, samples=100, seed=None)
    gaussian(img_in, sigma=2.0, nb_stddev=5)
    sobel(img_in)
    canny(img_in)

(upstream src/pillowfight/__init__.py -- not in the mount).
`img_in` is a PIL image (converted to RGBA, a PIL RGBA image is returned, as
upstream does) or an (H, W, 4) uint8 numpy array (an array is returned).

All work is done by libpillowfight_b200.so through its C-ABI
(include/pillowfight_b200.h); there is no CPU fallback and importing this
package fails if the library has not been built.
"""
import ctypes
import time

import numpy as np

from . import _capi
from ._capi import PillowfightB200Error, PfBitmap, lib  # noqa: F401

__all__ = ["ace", "gaussian", "sobel", "canny", "PillowfightB200Error",
           "device_count"]

_libc = ctypes.CDLL(None)


def device_count():
    return lib.pf_b200_device_count()


def _pil_pixels(img):
    """The pixels of a PIL image as an (H, W, 4) uint8 array in pinned memory.

    Upstream converts with img.convert("RGBA") and hands Pillow's bytes to C.
    Here RGBA and RGB images are packed band by band by Pillow's own raw
    encoder (what tobytes() uses) straight into a pooled pinned buffer: no
    35 MB bytes object per call (whose fresh pages cost more than the copy)
    and, for RGB, no conversion pass -- Pillow stores RGB as four bytes per
    pixel and none of the four operations reads the fourth.  Other modes are
    converted first, as upstream does."""
    from PIL import Image
    if img.mode not in ("RGBA", "RGB"):
        img = img.convert("RGBA")
    img.load()
    w, h = img.size
    arr = _pool.empty((h, w, 4))
    if w == 0 or h == 0:
        return arr
    rawmode = "RGBA" if img.mode == "RGBA" else "RGBX"
    flat = arr.reshape(-1)

    def pack(y_begin, y_end):
        rows = max(1, (4 << 20) // (4 * w))
        for y0 in range(y_begin, y_end, rows):
            y1 = min(y_end, y0 + rows)
            enc = Image._getencoder(img.mode, "raw", rawmode)
            enc.setimage(img.im, (0, y0, w, y1))
            off, end = 4 * w * y0, 4 * w * y1
            while True:
                _, err, data = enc.encode(end - off + 64)
                flat[off:off + len(data)] = np.frombuffer(data, dtype=np.uint8)
                off += len(data)
                if err:
                    break
            if err < 0 or off != end:
                raise RuntimeError("Pillow's raw encoder failed (%d)" % err)

    if arr.nbytes < (8 << 20):
        pack(0, h)
    else:
        # the copies run without the GIL: four threads roughly halve the time
        global _pack_pool
        if _pack_pool is None:
            from concurrent.futures import ThreadPoolExecutor
            _pack_pool = ThreadPoolExecutor(4, thread_name_prefix="pf_b200_pack")
        bands = [(h * i // 4, h * (i + 1) // 4) for i in range(4)]
        list(_pack_pool.map(lambda b: pack(*b), bands))
    return arr


_pack_pool = None


def _to_array(img_in):
    """-> (array (H, W, 4) uint8 C-contiguous, was_pil)"""
    if isinstance(img_in, np.ndarray):
        arr = np.ascontiguousarray(img_in, dtype=np.uint8)
        if arr.ndim != 3 or arr.shape[2] != 4:
            raise ValueError("expected an (H, W, 4) uint8 RGBA array")
        return arr, False
    return _pil_pixels(img_in), True


def _from_array(arr, was_pil):
    if not was_pil:
        return arr
    from PIL import Image
    # The image wraps the result buffer (pinned memory the device wrote by
    # DMA) instead of copying it twice; it keeps the buffer alive, and Pillow
    # copies on the first in-place modification (the image is read-only).
    return Image.frombuffer("RGBA", (arr.shape[1], arr.shape[0]), arr, "raw", "RGBA", 0, 1)


class _PinnedPool:
    """Result arrays live in pinned host memory (pf_b200_host_alloc): the
    device writes them by DMA directly, with no staging copy and no page
    faults on a fresh buffer.  A buffer returns to the pool when the array
    (and every view of it) has been collected; a few are kept per size."""
    KEEP = 4

    def __init__(self):
        self._free = {}

    def empty(self, shape):
        import weakref
        n = int(np.prod(shape))
        if n == 0:
            return np.empty(shape, dtype=np.uint8)
        bucket = self._free.setdefault(n, [])
        ptr = bucket.pop() if bucket else lib.pf_b200_host_alloc(n)
        if not ptr:
            return np.empty(shape, dtype=np.uint8)     # no device / no pinned memory: plain array
        raw = (ctypes.c_uint8 * n).from_address(ptr)
        weakref.finalize(raw, self._release, n, ptr)
        return np.frombuffer(raw, dtype=np.uint8).reshape(shape)

    def empty_like(self, arr):
        return self.empty(arr.shape)

    def _release(self, n, ptr):
        bucket = self._free.setdefault(n, [])
        if len(bucket) < self.KEEP:
            bucket.append(ptr)
        else:
            lib.pf_b200_host_free(ptr)


_pool = _PinnedPool()


def _bitmap(arr):
    bm = PfBitmap()
    bm.size.x = arr.shape[1]
    bm.size.y = arr.shape[0]
    bm.pixels = arr.ctypes.data
    return bm


def _run(fn, img_in, *args):
    arr, was_pil = _to_array(img_in)
    out = _pool.empty_like(arr)
    bi, bo = _bitmap(arr), _bitmap(out)
    _capi.check(fn(ctypes.byref(bi), ctypes.byref(bo), *args))
    return _from_array(out, was_pil)


def gaussian(img_in, sigma=2.0, nb_stddev=5):
    """Gaussian blur of R, G and B (upstream pillowfight.gaussian)."""
    return _run(lib.pf_b200_gaussian, img_in, float(sigma), int(nb_stddev))


def sobel(img_in):
    """Sobel gradient intensity (upstream pillowfight.sobel)."""
    return _run(lib.pf_b200_sobel, img_in)


def canny(img_in):
    """Canny edge map, 0 / 255 (upstream pillowfight.canny)."""
    return _run(lib.pf_b200_canny, img_in)


def ace_samples(samples, width, height, seed):
    """The sample table pf_apply_ace would draw after srand(seed)."""
    _libc.srand(ctypes.c_uint(int(seed) & 0xFFFFFFFF))
    xy = np.zeros(2 * int(samples), dtype=np.int32)
    lib.pf_b200_ace_draw_samples(int(samples), int(width), int(height),
                                 xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int)))
    return xy


def ace(img_in, slope=10, limit=1000, samples=100, seed=None):
    """Automatic Color Equalization (upstream pillowfight.ace).  The sample
    coordinates come from libc rand() seeded with `seed`, on the host."""
    if seed is None:
        seed = int(time.time())
    arr, was_pil = _to_array(img_in)
    xy = ace_samples(samples, arr.shape[1], arr.shape[0], seed)
    out = _pool.empty_like(arr)
    bi, bo = _bitmap(arr), _bitmap(out)
    _capi.check(lib.pf_b200_ace(ctypes.byref(bi), ctypes.byref(bo), int(samples),
                                xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int)),
                                float(slope), float(limit)))
    return _from_array(out, was_pil)
