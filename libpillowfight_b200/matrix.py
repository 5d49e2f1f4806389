"""numpy front end of the pf_dbl_matrix-level entry points
(pf_dbl_matrix_convolution, pf_gaussian_on_matrix, pf_sobel_on_matrix).
Matrices are float64 arrays of shape (height, width)."""
import ctypes

import numpy as np

from . import _capi
from ._capi import PfMatrix, lib


def _wrap(arr):
    arr = np.ascontiguousarray(arr, dtype=np.float64)
    if arr.ndim != 2:
        raise ValueError("expected a 2-D float64 matrix")
    m = PfMatrix()
    m.size.x, m.size.y = arr.shape[1], arr.shape[0]
    m.values = arr.ctypes.data_as(ctypes.POINTER(ctypes.c_double))
    return arr, m


def _take(m):
    """Copy a matrix the library allocated into numpy and release it."""
    if not m.values:
        code, msg = _capi.last_error()
        raise _capi.PillowfightB200Error(code, msg)
    out = np.ctypeslib.as_array(m.values, shape=(m.size.y, m.size.x)).copy()
    lib.pf_dbl_matrix_free(ctypes.byref(m))
    return out


def convolution(img, kernel):
    a, ma = _wrap(img)
    k, mk = _wrap(kernel)
    return _take(lib.pf_dbl_matrix_convolution(ctypes.byref(ma), ctypes.byref(mk)))


def gaussian_on_matrix(img, sigma=2.0, nb_stddev=5):
    a, ma = _wrap(img)
    return _take(lib.pf_gaussian_on_matrix(ctypes.byref(ma), float(sigma), int(nb_stddev)))


def sobel_on_matrix(img, kernel_x=None, kernel_y=None, gaussian_sigma=0.0, gaussian_stddev=0):
    """-> (g_x, g_y, intensity, direction)"""
    a, ma = _wrap(img)
    if kernel_x is None:
        mkx = PfMatrix.in_dll(lib, "g_pf_kernel_sobel_x")
    else:
        kx, mkx = _wrap(kernel_x)
    if kernel_y is None:
        mky = PfMatrix.in_dll(lib, "g_pf_kernel_sobel_y")
    else:
        ky, mky = _wrap(kernel_y)
    g = lib.pf_sobel_on_matrix(ctypes.byref(ma), ctypes.byref(mkx), ctypes.byref(mky),
                               float(gaussian_sigma), int(gaussian_stddev))
    if not g.g_x.values:
        code, msg = _capi.last_error()
        raise _capi.PillowfightB200Error(code, msg)
    return tuple(_take(m) for m in (g.g_x, g.g_y, g.intensity, g.direction))
