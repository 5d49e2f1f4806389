"""The C-ABI library loads and exports every function include/*.h declares.
No compute call is made here (no GPU needed)."""
import ctypes
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pillowfight_b200.h")


def declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"\b(pf_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


def test_header_declares_the_drop_in_entry_points():
    names = declared_functions()
    for required in ("pf_apply_ace", "pf_apply_gaussian", "pf_apply_sobel", "pf_apply_canny"):
        assert required in names


def test_library_exports_every_declared_symbol(pf):
    from libpillowfight_b200 import _capi
    raw = ctypes.CDLL(_capi.SO_PATH)
    names = declared_functions()
    assert len(names) >= 25
    for n in names:
        assert hasattr(raw, n), "libpillowfight_b200.so does not export %s" % n
        assert n in _capi.SIGNATURES, "%s has no ctypes signature in _capi.py" % n


def test_library_has_no_python_or_torch_dependency(pf):
    import subprocess
    from libpillowfight_b200 import _capi
    out = subprocess.run(["ldd", _capi.SO_PATH], capture_output=True, text=True).stdout
    assert "torch" not in out and "python" not in out


def test_struct_layout_matches_the_header(pf):
    from libpillowfight_b200 import _capi
    # struct pf_bitmap { struct { int x; int y; } size; union pf_pixel *pixels; }
    assert ctypes.sizeof(_capi.PfBitmap) == 16
    assert _capi.PfBitmap.pixels.offset == 8
    # union pf_pixel: r, g, b, a bytes in memory order
    px = np.array([[[1, 2, 3, 4]]], dtype=np.uint8)
    assert px.view(np.uint32)[0, 0, 0] == 0x04030201


def test_sample_draw_matches_the_oracle(pf, oracle):
    # host-only code path: libc rand(), x then y
    a = pf.ace_samples(100, 2480, 3508, 12345)
    b = oracle.ace_draw_samples(100, 2480, 3508, 12345)
    assert np.array_equal(a, b)
    assert (a[0::2] < 2480).all() and (a[1::2] < 3508).all()


def test_fails_loudly_without_a_device(pf):
    if pf.device_count() > 0:
        import pytest
        pytest.skip("a CUDA device is present")
    img = np.zeros((8, 8, 4), dtype=np.uint8)
    for call in (lambda: pf.gaussian(img), lambda: pf.sobel(img), lambda: pf.canny(img),
                 lambda: pf.ace(img, seed=1)):
        try:
            call()
        except pf.PillowfightB200Error as e:
            assert e.code == 2 and "no CPU path" in str(e)
        else:
            raise AssertionError("a call succeeded without a CUDA device")


def test_argument_errors_are_reported(pf):
    from libpillowfight_b200 import _capi
    bi = _capi.PfBitmap()
    bo = _capi.PfBitmap()
    rc = _capi.lib.pf_b200_sobel(ctypes.byref(bi), ctypes.byref(bo))
    assert rc == _capi.PF_B200_EINVAL
    code, msg = _capi.last_error()
    assert code == _capi.PF_B200_EINVAL and "NULL" in msg


def test_reference_module_name_and_signatures(pf):
    """`import pillowfight` exposes the four functions with upstream's
    parameter names and defaults."""
    import inspect
    import pillowfight
    want = {
        "ace": [("img_in", inspect.Parameter.empty), ("slope", 10), ("limit", 1000),
                ("samples", 100), ("seed", None)],
        "gaussian": [("img_in", inspect.Parameter.empty), ("sigma", 2.0), ("nb_stddev", 5)],
        "sobel": [("img_in", inspect.Parameter.empty)],
        "canny": [("img_in", inspect.Parameter.empty)],
    }
    for name, params in want.items():
        sig = inspect.signature(getattr(pillowfight, name))
        got = [(p.name, p.default) for p in sig.parameters.values()]
        assert got == params, (name, got)
