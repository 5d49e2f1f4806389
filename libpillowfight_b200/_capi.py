"""ctypes binding of libpillowfight_b200.so (include/pillowfight_b200.h).

The shared library is the product; this file only declares its C-ABI to
Python.  There is no fallback of any kind: if the library has not been built
(python libpillowfight_b200/build.py) importing this module raises.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# PF_B200_LIBRARY: another build of the same library (kernel experiments)
SO_PATH = os.environ.get("PF_B200_LIBRARY") or os.path.join(_HERE, "libpillowfight_b200.so")

PF_B200_OK = 0
PF_B200_EINVAL = 1
PF_B200_ENODEV = 2
PF_B200_ECUDA = 3
PF_B200_ETOOBIG = 4


class PillowfightB200Error(RuntimeError):
    def __init__(self, code, message):
        super().__init__("libpillowfight_b200 error %d: %s" % (code, message))
        self.code = code


class _Size(ctypes.Structure):
    _fields_ = [("x", ctypes.c_int), ("y", ctypes.c_int)]


class PfBitmap(ctypes.Structure):
    """struct pf_bitmap"""
    _fields_ = [("size", _Size), ("pixels", ctypes.c_void_p)]


class PfMatrix(ctypes.Structure):
    """struct pf_dbl_matrix"""
    _fields_ = [("size", _Size), ("values", ctypes.POINTER(ctypes.c_double))]


class PfGradients(ctypes.Structure):
    """struct pf_gradient_matrixes"""
    _fields_ = [("g_x", PfMatrix), ("g_y", PfMatrix), ("intensity", PfMatrix),
                ("direction", PfMatrix)]


if not os.path.exists(SO_PATH):
    raise ImportError(
        "libpillowfight_b200.so is not built (expected at %s). Build it with "
        "`python libpillowfight_b200/build.py`; there is no CPU fallback." % SO_PATH)

lib = ctypes.CDLL(SO_PATH)

_bp = ctypes.POINTER(PfBitmap)
_u32p = ctypes.c_void_p      # device pointers travel as integers
_vp = ctypes.c_void_p
_ip = ctypes.POINTER(ctypes.c_int)

# every symbol include/pillowfight_b200.h declares: (restype, argtypes)
SIGNATURES = {
    "pf_apply_ace": (None, [_bp, _bp, ctypes.c_int, ctypes.c_double, ctypes.c_double, ctypes.c_int]),
    "pf_apply_gaussian": (None, [_bp, _bp, ctypes.c_double, ctypes.c_int]),
    "pf_apply_sobel": (None, [_bp, _bp]),
    "pf_apply_canny": (None, [_bp, _bp]),
    "pf_ace": (None, [_bp, _bp, ctypes.c_int, ctypes.c_double, ctypes.c_double, ctypes.c_int]),
    "pf_gaussian": (None, [_bp, _bp, ctypes.c_double, ctypes.c_int]),
    "pf_sobel": (None, [_bp, _bp]),
    "pf_canny": (None, [_bp, _bp]),
    "pf_dbl_matrix_new": (PfMatrix, [ctypes.c_int, ctypes.c_int]),
    "pf_dbl_matrix_free": (None, [ctypes.POINTER(PfMatrix)]),
    "pf_dbl_matrix_convolution": (PfMatrix, [ctypes.POINTER(PfMatrix), ctypes.POINTER(PfMatrix)]),
    "pf_gaussian_on_matrix": (PfMatrix, [ctypes.POINTER(PfMatrix), ctypes.c_double, ctypes.c_int]),
    "pf_sobel_on_matrix": (PfGradients, [ctypes.POINTER(PfMatrix), ctypes.POINTER(PfMatrix),
                                         ctypes.POINTER(PfMatrix), ctypes.c_double, ctypes.c_int]),
    "pf_b200_last_error": (ctypes.c_int, []),
    "pf_b200_last_error_string": (ctypes.c_char_p, []),
    "pf_b200_device_count": (ctypes.c_int, []),
    "pf_b200_ace": (ctypes.c_int, [_bp, _bp, ctypes.c_int, _ip, ctypes.c_double, ctypes.c_double]),
    "pf_b200_gaussian": (ctypes.c_int, [_bp, _bp, ctypes.c_double, ctypes.c_int]),
    "pf_b200_sobel": (ctypes.c_int, [_bp, _bp]),
    "pf_b200_canny": (ctypes.c_int, [_bp, _bp]),
    "pf_b200_ace_multi_gpu": (ctypes.c_int, [_bp, _bp, ctypes.c_int, _ip, ctypes.c_double, ctypes.c_double,
                                             _ip, ctypes.c_int]),
    "pf_b200_page_step": (ctypes.c_int, [_bp, _bp, _bp, _bp, _bp, ctypes.c_int, _ip, ctypes.c_double,
                                         ctypes.c_double, ctypes.c_double, ctypes.c_int]),
    "pf_b200_ace_draw_samples": (None, [ctypes.c_int, ctypes.c_int, ctypes.c_int, _ip]),
    "pf_b200_host_alloc": (ctypes.c_void_p, [ctypes.c_size_t]),
    "pf_b200_host_free": (None, [ctypes.c_void_p]),
    "pf_b200_dev_gaussian": (ctypes.c_int, [_u32p, _u32p, ctypes.c_int, ctypes.c_int,
                                            ctypes.c_double, ctypes.c_int, _vp]),
    "pf_b200_dev_sobel": (ctypes.c_int, [_u32p, _u32p, ctypes.c_int, ctypes.c_int, _vp]),
    "pf_b200_canny_scratch_bytes": (ctypes.c_size_t, [ctypes.c_int, ctypes.c_int]),
    "pf_b200_dev_canny": (ctypes.c_int, [_u32p, _u32p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                         _vp, ctypes.c_size_t, _vp]),
    "pf_b200_dev_canny_noalloc": (ctypes.c_int, [_u32p, _u32p, ctypes.c_int, ctypes.c_int, _vp]),
    "pf_b200_dev_gaussian_batch": (ctypes.c_int, [_u32p, _u32p, ctypes.c_int, ctypes.c_int,
                                                  ctypes.c_int, ctypes.c_double, ctypes.c_int, _vp]),
    "pf_b200_dev_sobel_batch": (ctypes.c_int, [_u32p, _u32p, ctypes.c_int, ctypes.c_int,
                                               ctypes.c_int, _vp]),
    "pf_b200_ace_minmax_bytes": (ctypes.c_size_t, [ctypes.c_int, ctypes.c_int]),
    "pf_b200_ace_scores_bytes": (ctypes.c_size_t, [ctypes.c_int, ctypes.c_int]),
    "pf_b200_dev_ace_minmax_init": (ctypes.c_int, [_vp, ctypes.c_int, ctypes.c_int, _vp]),
    # (d_in, w, h, row_begin, row_end, d_samples, n, slope, limit, d_scores, d_minmax[, d_out], stream)
    "pf_b200_dev_ace_scores": (ctypes.c_int, [_u32p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                              ctypes.c_int, _vp, ctypes.c_int, ctypes.c_double,
                                              ctypes.c_double, _vp, _vp, _vp]),
    "pf_b200_dev_ace_refine": (ctypes.c_int, [_u32p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                              ctypes.c_int, _vp, ctypes.c_int, ctypes.c_double,
                                              ctypes.c_double, _vp, _vp, _vp]),
    "pf_b200_dev_ace_scale": (ctypes.c_int, [_u32p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                             ctypes.c_int, _vp, ctypes.c_int, ctypes.c_double,
                                             ctypes.c_double, _vp, _vp, _u32p, _vp]),
    # (d_band, w, h, row_begin, row_end, d_samples, d_sample_px, n, slope, limit, d_scores, d_minmax, ...)
    "pf_b200_dev_ace_scores_band": (ctypes.c_int, [_u32p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                                   ctypes.c_int, _vp, _vp, ctypes.c_int, ctypes.c_double,
                                                   ctypes.c_double, _vp, _vp, _vp]),
    "pf_b200_dev_ace_refine_band": (ctypes.c_int, [_u32p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                                   ctypes.c_int, _vp, _vp, ctypes.c_int, ctypes.c_double,
                                                   ctypes.c_double, _vp, _vp, _vp]),
    "pf_b200_dev_ace_scale_band": (ctypes.c_int, [_u32p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                                  ctypes.c_int, _vp, _vp, ctypes.c_int, ctypes.c_double,
                                                  ctypes.c_double, _vp, _vp, _u32p, _vp]),
    "pf_b200_dev_ace_gather_samples": (ctypes.c_int, [_u32p, ctypes.c_int, ctypes.c_int, _vp, ctypes.c_int,
                                                      _vp, _vp]),
    "pf_b200_dev_ace_scale_to": (ctypes.c_int, [_u32p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                                ctypes.c_int, _vp, _vp, ctypes.c_int, ctypes.c_double,
                                                ctypes.c_double, _vp, _vp,
                                                ctypes.POINTER(ctypes.c_void_p), ctypes.c_int, _vp]),
    "pf_b200_dev_ace_scale_multicast": (ctypes.c_int, [_u32p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                                       ctypes.c_int, _vp, _vp, ctypes.c_int, ctypes.c_double,
                                                       ctypes.c_double, _vp, _vp, _vp, _vp]),
    "pf_b200_ace_workspace_bytes": (ctypes.c_size_t, [ctypes.c_int, ctypes.c_int]),
    "pf_b200_dev_ace": (ctypes.c_int, [_u32p, _u32p, ctypes.c_int, ctypes.c_int, _vp, ctypes.c_int,
                                       ctypes.c_double, ctypes.c_double, _vp, ctypes.c_size_t, _vp]),
    "pf_b200_fp64_peak_tflops": (ctypes.c_double, []),
    "pf_b200_fp32_peak_tflops": (ctypes.c_double, []),
    "pf_b200_selftest_rsqrt_error": (ctypes.c_double, [ctypes.c_int]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)      # AttributeError here = symbol missing from the .so
    _fn.restype = _res
    _fn.argtypes = _args


def last_error():
    msg = lib.pf_b200_last_error_string()
    return lib.pf_b200_last_error(), (msg.decode("utf-8", "replace") if msg else "")


def check(rc):
    """Raise PillowfightB200Error for a non-zero status of a pf_b200_* call."""
    if rc != PF_B200_OK:
        code, msg = last_error()
        raise PillowfightB200Error(rc, msg or "status %d" % code)
