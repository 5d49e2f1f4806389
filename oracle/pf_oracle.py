"""ctypes front end of the CPU oracle (oracle/pf_oracle.c).

Test infrastructure only: imported by tests/, __graft_entry__.smoke() and
bench.py's CPU-baseline legs, never by libpillowfight_b200/.  PARITY UNPINNED:
see the header of pf_oracle.c and DESIGN.md section 0.

Images are numpy uint8 arrays of shape (height, width, 4), RGBA, C-contiguous.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libpf_oracle.so")


class _Size(ctypes.Structure):
    _fields_ = [("x", ctypes.c_int), ("y", ctypes.c_int)]


class Bitmap(ctypes.Structure):
    _fields_ = [("size", _Size), ("pixels", ctypes.c_void_p)]


class Matrix(ctypes.Structure):
    _fields_ = [("size", _Size), ("values", ctypes.POINTER(ctypes.c_double))]


class Gradients(ctypes.Structure):
    _fields_ = [("g_x", Matrix), ("g_y", Matrix), ("intensity", Matrix), ("direction", Matrix)]


def build(force=False):
    """Compile libpf_oracle.so with oracle/Makefile (gcc)."""
    src = os.path.join(_HERE, "pf_oracle.c")
    if (force or not os.path.exists(_SO)
            or os.path.getmtime(_SO) < os.path.getmtime(src)):
        subprocess.check_call(["make", "-s", "-C", _HERE])
    return _SO


_lib = None
_libc = ctypes.CDLL(None)


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        _lib.pfo_gaussian_taps.argtypes = [ctypes.c_double, ctypes.c_int,
                                           ctypes.POINTER(ctypes.c_double)]
        _lib.pfo_canny_thresholds.argtypes = [ctypes.POINTER(ctypes.c_double)] * 2
        bp = ctypes.POINTER(Bitmap)
        _lib.pfo_gaussian.argtypes = [bp, bp, ctypes.c_double, ctypes.c_int]
        _lib.pfo_sobel.argtypes = [bp, bp]
        _lib.pfo_canny.argtypes = [bp, bp]
        _lib.pfo_ace.argtypes = [bp, bp, ctypes.c_int, ctypes.c_double,
                                 ctypes.c_double, ctypes.c_int]
        ip = ctypes.POINTER(ctypes.c_int)
        _lib.pfo_ace_with_samples.argtypes = [bp, bp, ctypes.c_int, ip,
                                              ctypes.c_double, ctypes.c_double,
                                              ctypes.c_int]
        _lib.pfo_ace_draw_samples.argtypes = [ctypes.c_int, ctypes.c_int,
                                              ctypes.c_int, ip]
        _lib.pfo_ace_adjust_band.argtypes = [bp, ctypes.c_int, ip,
                                             ctypes.c_double, ctypes.c_double,
                                             ctypes.c_int, ctypes.c_int,
                                             ctypes.c_int,
                                             ctypes.POINTER(ctypes.c_double)]
        _lib.pfo_gaussian_on_matrix.argtypes = [ctypes.POINTER(Matrix),
                                                ctypes.c_double, ctypes.c_int]
        _lib.pfo_gaussian_on_matrix.restype = Matrix
        _lib.pfo_matrix_free.argtypes = [ctypes.POINTER(Matrix)]
        _lib.pfo_convolution.argtypes = [ctypes.POINTER(Matrix), ctypes.POINTER(Matrix)]
        _lib.pfo_convolution.restype = Matrix
        _lib.pfo_sobel_on_matrix.argtypes = [ctypes.POINTER(Matrix)]
        _lib.pfo_sobel_on_matrix.restype = Gradients
    return _lib


def _check(img):
    img = np.ascontiguousarray(img, dtype=np.uint8)
    if img.ndim != 3 or img.shape[2] != 4:
        raise ValueError("expected an (H, W, 4) uint8 RGBA array")
    return img


def _bitmap(arr):
    return Bitmap(_Size(arr.shape[1], arr.shape[0]), arr.ctypes.data)


def _unary(fn, img, *args):
    img = _check(img)
    out = np.empty_like(img)
    bi, bo = _bitmap(img), _bitmap(out)
    fn(ctypes.byref(bi), ctypes.byref(bo), *args)
    return out


def srand(seed):
    """Seed libc rand() -- the RNG the ACE sample table is drawn from."""
    _libc.srand(ctypes.c_uint(seed & 0xFFFFFFFF))


def gaussian_taps(sigma, nb_stddev):
    taps = np.zeros(nb_stddev, dtype=np.float64)
    lib().pfo_gaussian_taps(sigma, nb_stddev,
                            taps.ctypes.data_as(ctypes.POINTER(ctypes.c_double)))
    return taps


def canny_thresholds():
    lo, hi = ctypes.c_double(), ctypes.c_double()
    lib().pfo_canny_thresholds(ctypes.byref(lo), ctypes.byref(hi))
    return lo.value, hi.value


def gaussian(img, sigma=2.0, nb_stddev=5):
    return _unary(lib().pfo_gaussian, img, sigma, nb_stddev)


def gaussian_on_matrix(mat, sigma=2.0, nb_stddev=5):
    mat = np.ascontiguousarray(mat, dtype=np.float64)
    m = Matrix(_Size(mat.shape[1], mat.shape[0]),
               mat.ctypes.data_as(ctypes.POINTER(ctypes.c_double)))
    r = lib().pfo_gaussian_on_matrix(ctypes.byref(m), sigma, nb_stddev)
    out = np.ctypeslib.as_array(r.values, shape=mat.shape).copy()
    lib().pfo_matrix_free(ctypes.byref(r))
    return out


def _matrix(mat):
    mat = np.ascontiguousarray(mat, dtype=np.float64)
    return mat, Matrix(_Size(mat.shape[1], mat.shape[0]),
                       mat.ctypes.data_as(ctypes.POINTER(ctypes.c_double)))


def _take(r):
    out = np.ctypeslib.as_array(r.values, shape=(r.size.y, r.size.x)).copy()
    lib().pfo_matrix_free(ctypes.byref(r))
    return out


def convolution(mat, kernel):
    a, ma = _matrix(mat)
    k, mk = _matrix(kernel)
    return _take(lib().pfo_convolution(ctypes.byref(ma), ctypes.byref(mk)))


def sobel_on_matrix(mat):
    """-> (g_x, g_y, intensity, direction) with the reference's 3x3 kernels"""
    a, ma = _matrix(mat)
    g = lib().pfo_sobel_on_matrix(ctypes.byref(ma))
    return tuple(_take(m) for m in (g.g_x, g.g_y, g.intensity, g.direction))


def sobel(img):
    return _unary(lib().pfo_sobel, img)


def canny(img):
    return _unary(lib().pfo_canny, img)


def ace_draw_samples(nb_samples, width, height, seed):
    srand(seed)
    xy = np.zeros(2 * nb_samples, dtype=np.int32)
    lib().pfo_ace_draw_samples(nb_samples, width, height,
                               xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int)))
    return xy


def ace_with_samples(img, xy, slope=10.0, limit=1000.0, nb_threads=1):
    xy = np.ascontiguousarray(xy, dtype=np.int32)
    return _unary(lib().pfo_ace_with_samples, img, len(xy) // 2,
                  xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int)),
                  slope, limit, nb_threads)


def ace(img, slope=10.0, limit=1000.0, samples=100, seed=0, nb_threads=1):
    srand(seed)
    return _unary(lib().pfo_ace, img, samples, slope, limit, nb_threads)


def ace_scores(img, xy, slope=10.0, limit=1000.0, row_begin=0, row_end=None,
               nb_threads=1):
    """Adjustment pass only: returns the (3, H, W) float64 score planes (rows
    outside [row_begin, row_end) are left at 0)."""
    img = _check(img)
    h, w = img.shape[:2]
    if row_end is None:
        row_end = h
    xy = np.ascontiguousarray(xy, dtype=np.int32)
    scores = np.zeros((3, h, w), dtype=np.float64)
    bi = _bitmap(img)
    lib().pfo_ace_adjust_band(ctypes.byref(bi), len(xy) // 2,
                              xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int)),
                              slope, limit, row_begin, row_end, nb_threads,
                              scores.ctypes.data_as(ctypes.POINTER(ctypes.c_double)))
    return scores
