"""pf_dbl_matrix-level entry points (pf_dbl_matrix_convolution,
pf_gaussian_on_matrix, pf_sobel_on_matrix) against the oracle.

Convolutions follow the reference's operation order with separate multiply and
add roundings: bit-exact.  hypot / atan2 come from CUDA's libm, the oracle's
from glibc: both are accurate to about an ulp, so intensity and direction are
held to 4 ulp."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ULP = np.finfo(np.float64).eps


@pytest.fixture(scope="module")
def mx(pf):
    if pf.device_count() <= 0:
        pytest.fail("no CUDA device: the gpu-marked tests must run on the B200 box")
    from libpillowfight_b200 import matrix
    return matrix


@pytest.mark.parametrize("w,h", [(1, 1), (7, 5), (64, 33), (301, 200)])
def test_convolution_bit_exact(mx, oracle, w, h):
    rng = np.random.default_rng(w * 100 + h)
    m = rng.normal(size=(h, w)) * 100.0
    for k in (rng.normal(size=(3, 3)), rng.normal(size=(1, 5)), rng.normal(size=(5, 1)),
              rng.normal(size=(2, 4)), np.array([[2.5]])):
        assert np.array_equal(mx.convolution(m, k), oracle.convolution(m, k))


@pytest.mark.parametrize("sigma,n", [(2.0, 5), (0.0, 5), (1.0, 3), (3.0, 9), (2.0, 4), (5.0, 21)])
def test_gaussian_on_matrix_bit_exact(mx, oracle, sigma, n):
    rng = np.random.default_rng(n)
    m = rng.uniform(0.0, 255.0, size=(150, 211))
    assert np.array_equal(mx.gaussian_on_matrix(m, sigma, n), oracle.gaussian_on_matrix(m, sigma, n))


def test_sobel_on_matrix(mx, oracle):
    from libpillowfight_b200.pages import synthetic_page
    page = synthetic_page(320, 240, 3)
    gray = (page[..., :3].astype(np.int32).sum(axis=-1) // 3).astype(np.float64)
    gx, gy, inten, direc = mx.sobel_on_matrix(gray)
    rgx, rgy, rint, rdir = oracle.sobel_on_matrix(gray)
    assert np.array_equal(gx, rgx) and np.array_equal(gy, rgy)
    assert np.all(np.abs(inten - rint) <= 4 * ULP * np.maximum(np.abs(rint), 1e-300))
    assert np.all(np.abs(direc - rdir) <= 4 * ULP * np.pi)
    # non-integer input and explicit kernels
    rng = np.random.default_rng(1)
    m = rng.normal(size=(90, 120))
    kx = np.array([[-1.0, 0.0, 1.0], [-2.0, 0.0, 2.0], [-1.0, 0.0, 1.0]])
    gx2, gy2, i2, d2 = mx.sobel_on_matrix(m, kx, kx.T.copy())
    r = oracle.sobel_on_matrix(m)
    assert np.array_equal(gx2, r[0]) and np.array_equal(gy2, r[1])
    assert np.all(np.abs(i2 - r[2]) <= 4 * ULP * np.abs(r[2]) + 1e-300)


def test_sobel_on_matrix_with_gradient_blur(mx, oracle):
    rng = np.random.default_rng(2)
    m = rng.uniform(0, 255, size=(64, 80))
    gx, gy, inten, direc = mx.sobel_on_matrix(m, gaussian_sigma=1.5, gaussian_stddev=5)
    r = oracle.sobel_on_matrix(m)
    rgx = oracle.gaussian_on_matrix(r[0], 1.5, 5)
    rgy = oracle.gaussian_on_matrix(r[1], 1.5, 5)
    assert np.array_equal(gx, rgx) and np.array_equal(gy, rgy)
    assert np.all(np.abs(inten - np.hypot(rgx, rgy)) <= 4 * ULP * np.hypot(rgx, rgy) + 1e-300)


def test_matrix_errors(mx):
    from libpillowfight_b200 import _capi
    with pytest.raises(_capi.PillowfightB200Error):
        mx.convolution(np.zeros((4, 4)), np.zeros((9, 9)))      # more than 64 taps
    with pytest.raises(_capi.PillowfightB200Error):
        mx.gaussian_on_matrix(np.zeros((4, 4)), 2.0, 0)
