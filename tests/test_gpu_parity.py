"""Parity of the CUDA path with the CPU oracle, through the C-ABI.

Bar (BASELINE.json north_star): bit-exact for the integer stages (Sobel
intensity, Canny edge maps); for the floating-point stages at most 0.1 % of
channel values may differ, and then by no more than 1 LSB.  This build holds
the floating-point stages to bit-exactness too: the Gaussian falls back to the
reference's operation order wherever its fast sums could round differently,
and ACE recomputes in the reference's arithmetic every pixel its bounded FP32
pass cannot decide.
"""
import ctypes

import numpy as np
import pytest

from libpillowfight_b200.pages import A4_300DPI, synthetic_page, uniform_random_page

pytestmark = pytest.mark.gpu

FP_MAX_FRACTION = 1e-3     # <= 0.1 % of channel values ...
FP_MAX_LSB = 1             # ... off by at most 1


def assert_fp_parity(got, ref, what):
    """north_star's bar for the floating-point stages (kept for reference)."""
    assert got.shape == ref.shape
    d = np.abs(got.astype(np.int16) - ref.astype(np.int16))
    assert d.max() <= FP_MAX_LSB, "%s: max diff %d" % (what, d.max())
    frac = (d[..., :3] != 0).mean()
    assert frac <= FP_MAX_FRACTION, "%s: %.5f %% of values differ" % (what, 100 * frac)
    return frac


def assert_exact(got, ref, what):
    """What ACE is held to here: every channel value equal to the oracle's."""
    assert got.shape == ref.shape
    if not np.array_equal(got, ref):
        d = np.abs(got.astype(np.int16) - ref.astype(np.int16))
        bad = np.argwhere(d != 0)
        raise AssertionError("%s: %d of %d values differ (max %d), first at %s: got %s, oracle %s" % (
            what, len(bad), d.size, d.max(), bad[0].tolist(), got[tuple(bad[0][:2])].tolist(),
            ref[tuple(bad[0][:2])].tolist()))


SIZES = [(1, 1), (2, 3), (5, 4), (7, 9), (33, 17), (64, 48), (130, 67), (257, 129),
         (256, 200), (620, 877),
         # widths that are multiples of 4 take the vectorised kernels: tiny and
         # one-warp-column / many-warp-column shapes
         (4, 1), (8, 3), (12, 2), (128, 1), (120, 7), (124, 19), (244, 5), (1000, 3)]


@pytest.fixture(scope="module")
def gpu(pf):
    if pf.device_count() <= 0:
        pytest.fail("no CUDA device: the gpu-marked tests must run on the B200 box")
    return pf


@pytest.mark.parametrize("w,h", SIZES)
def test_gaussian_bit_exact(gpu, oracle, w, h):
    page = uniform_random_page(w, h, seed=w * 1000 + h)
    assert np.array_equal(gpu.gaussian(page, 2.0, 5), oracle.gaussian(page, 2.0, 5))


@pytest.mark.parametrize("sigma,n", [(2.0, 5), (0.0, 5), (1.0, 3), (3.0, 7), (2.5, 9),
                                     (2.0, 4), (1.0, 1), (4.0, 11), (2.0, 6)])
def test_gaussian_tap_counts(gpu, oracle, sigma, n):
    page = synthetic_page(150, 90, 2)
    assert np.array_equal(gpu.gaussian(page, sigma, n), oracle.gaussian(page, sigma, n))


def test_gaussian_flat_regions(gpu, oracle):
    # saturated / flat areas are where a different rounding order would show
    # up as whole regions off by one
    for v in (255, 235, 200, 1, 0):
        page = np.full((40, 52, 4), v, dtype=np.uint8)
        page[..., 3] = 255
        assert np.array_equal(gpu.gaussian(page, 2.0, 5), oracle.gaussian(page, 2.0, 5))


@pytest.mark.parametrize("w,h", SIZES)
def test_sobel_bit_exact(gpu, oracle, w, h):
    page = uniform_random_page(w, h, seed=7 + w * 1000 + h)
    assert np.array_equal(gpu.sobel(page), oracle.sobel(page))


@pytest.mark.parametrize("w,h", SIZES)
def test_canny_bit_exact_random(gpu, oracle, w, h):
    page = uniform_random_page(w, h, seed=11 + w * 1000 + h)
    assert np.array_equal(gpu.canny(page), oracle.canny(page))


@pytest.mark.parametrize("idx", [0, 1, 2])
def test_canny_bit_exact_document(gpu, oracle, idx):
    page = synthetic_page(620, 877, idx)
    got, ref = gpu.canny(page), oracle.canny(page)
    assert np.array_equal(got, ref)
    assert (ref[..., 0] == 255).any()


def test_canny_adversarial_pages(gpu, oracle):
    flat = np.full((100, 120, 4), 128, dtype=np.uint8)
    flat[..., 3] = 255
    assert np.array_equal(gpu.canny(flat), oracle.canny(flat))
    yy, xx = np.mgrid[0:100, 0:120]
    cb = np.where((xx + yy) % 2 == 0, 255, 0).astype(np.uint8)
    page = np.stack([cb, cb, cb, np.full_like(cb, 255)], axis=-1)
    assert np.array_equal(gpu.canny(page), oracle.canny(page))


def test_canny_long_weak_chains(gpu, oracle):
    # A weak edge hundreds of pixels long, kept only because one end touches a
    # strong segment: the hysteresis has to follow the whole chain, across
    # thread, warp and strip boundaries.  A second, unconnected weak edge of
    # the same contrast must vanish.
    h, w = 700, 300
    g = np.zeros((h, w), dtype=np.uint8)
    g[:, 100:] = 50            # |g_x| = 200: weak (174.9 < 200 <= 349.9)
    g[:30, 100:] = 120         # |g_x| = 480: strong near the top
    g[60:640, 200:260] += 50   # a weak-contrast rectangle touching nothing strong
    page = np.stack([g, g, g, np.full_like(g, 255)], axis=-1)
    got, ref = gpu.canny(page), oracle.canny(page)
    assert np.array_equal(got, ref)
    assert ref[600, 99:101, 0].any()        # the chain reached the bottom
    assert not ref[100:600, 198:202, 0].any() and not ref[100:600, 258:262, 0].any()


@pytest.mark.parametrize("h", [60, 200, 420])
def test_canny_short_list_long_chain(gpu, oracle, h):
    # The same on a page small enough for the short-list path (one block, shared
    # memory): chains shorter and longer than the passes that spread "kept" along
    # the edges, so both that and the component labelling behind it decide pages.
    w = 120
    g = np.zeros((h, w), dtype=np.uint8)
    g[:, 40:] = 50
    g[:10, 40:] = 120
    g[20:h - 8, 80:100] += 50
    page = np.stack([g, g, g, np.full_like(g, 255)], axis=-1)
    got, ref = gpu.canny(page), oracle.canny(page)
    assert np.array_equal(got, ref)
    assert ref[h - 5, 39:41, 0].any()
    assert not ref[30:h - 16, 78:82, 0].any()


def test_canny_list_lengths_around_the_short_list_limit(gpu, oracle):
    """Work lists just under and just over the 2048 entries that one block
    resolves in shared memory: both paths, and the hand-over between them."""
    import torch
    from libpillowfight_b200 import _capi
    lib = _capi.lib
    dev = torch.device("cuda:0")
    seen = []
    for h in (580, 600, 612, 614, 615, 616, 618, 640):
        w = 120
        g = np.zeros((h, w), dtype=np.uint8)
        g[:, 40:] = 50                 # a weak vertical edge: three weak pixels per row
        g[:10, 40:] = 120              # strong at the top: the whole chain is kept
        page = np.stack([g, g, g, np.full_like(g, 255)], axis=-1)
        tin = torch.from_numpy(page).to(dev)
        tout = torch.empty_like(tin)
        scratch = torch.empty(lib.pf_b200_canny_scratch_bytes(w, h), dtype=torch.uint8, device=dev)
        _capi.check(lib.pf_b200_dev_canny(tin.data_ptr(), tout.data_ptr(), w, h, 1,
                                          scratch.data_ptr(), scratch.numel(), 0))
        torch.cuda.synchronize()
        seen.append(int(scratch[:16].view(torch.int32)[3].item()))     # length of the list just resolved
        assert np.array_equal(tout.cpu().numpy(), oracle.canny(page)), "h = %d, %d weak pixels" % (h, seen[-1])
    assert min(seen) < 2048 < max(seen), seen


def test_canny_snake_fills_queues(gpu, oracle):
    # dense weak texture seeded by a few strong pixels: a long work list, so the
    # grid-wide union-find path of the hysteresis kernel
    rng = np.random.default_rng(5)
    h, w = 256, 256
    g = (rng.integers(0, 2, size=(h, w)) * 45).astype(np.uint8)
    g[::64, ::64] = 255
    page = np.stack([g, g, g, np.full_like(g, 255)], axis=-1)
    assert np.array_equal(gpu.canny(page), oracle.canny(page))


@pytest.mark.parametrize("w,h,n", [(64, 48, 8), (200, 150, 30), (130, 257, 100), (620, 877, 100)])
def test_ace_parity(gpu, oracle, w, h, n):
    page = synthetic_page(w, h, 4)
    xy = oracle.ace_draw_samples(n, w, h, 12345)
    ref = oracle.ace_with_samples(page, xy, 10.0, 1000.0, nb_threads=8)
    got = gpu.ace(page, 10, 1000, n, seed=12345)     # draws the same table itself
    assert_exact(got, ref, "ace %dx%d" % (w, h))
    assert (got[..., 3] == 255).all()


@pytest.mark.parametrize("slope,limit", [(10.0, 1000.0), (5.0, 500.0), (7.5, 1000.0),
                                         (1.0, 300.0), (20.0, 100.0), (3.0, 1000.0),
                                         (0.1, 0.7), (1000.0, 999.0), (3.3, 841.5)])
def test_ace_slope_limit_variants(gpu, oracle, slope, limit):
    # integer and fractional limit/slope, saturation reached early, late, never
    page = uniform_random_page(160, 120, seed=9)
    xy = oracle.ace_draw_samples(40, 160, 120, 3)
    ref = oracle.ace_with_samples(page, xy, slope, limit, nb_threads=4)
    got = gpu.ace(page, slope, limit, 40, seed=3)
    assert_exact(got, ref, "ace slope=%g limit=%g" % (slope, limit))


def test_ace_pixels_without_any_sample(gpu, oracle):
    """One sample in the middle of the page: every pixel closer than height/5
    has no contributing sample at all (0/0 in the reference -> 0 in the output)."""
    from libpillowfight_b200 import _capi
    page = uniform_random_page(60, 50, seed=17)
    for xy in (np.array([30, 25], dtype=np.int32), np.array([0, 0, 59, 49], dtype=np.int32)):
        ref = oracle.ace_with_samples(page, xy, 10.0, 1000.0)
        out = np.empty_like(page)
        bi, bo = _capi.PfBitmap(), _capi.PfBitmap()
        bi.size.x, bi.size.y, bi.pixels = 60, 50, page.ctypes.data
        bo.size.x, bo.size.y, bo.pixels = 60, 50, out.ctypes.data
        _capi.check(_capi.lib.pf_b200_ace(ctypes.byref(bi), ctypes.byref(bo), len(xy) // 2,
                                          xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int)),
                                          10.0, 1000.0))
        # With one or two samples the weights cancel and every score is an
        # exact ratio c/100: many land exactly on a quantisation boundary,
        # where the last bit of the reference's own operation order decides --
        # those pixels are the ones the exact pass recomputes.
        assert_exact(out, ref, "ace with %d sample(s)" % (len(xy) // 2))
        if len(xy) == 2:
            near = ref[22:29, 27:34, :3]        # closer than 50/5 rows to the sample
            assert (near == 0).all() and (out[22:29, 27:34, :3] == 0).all()


def test_ace_flat_and_tied_pages(gpu, oracle):
    """Pages where many pixels score exactly the same in the reference: a
    difference of 0, +limit or -limit to every sample gives exactly 0, 1, -1.
    The rescale divides by (max - min): a flat channel is 0/0 -> 0 everywhere,
    and a tie AT the maximum must come out as 255 for every tied pixel."""
    w, h, n = 150, 90, 20
    xy = oracle.ace_draw_samples(n, w, h, 4)
    sample_px = set(zip(xy[0::2].tolist(), xy[1::2].tolist()))
    flat = np.full((h, w, 4), 255, dtype=np.uint8)
    flat[..., :3] = (200, 17, 255)
    mark = flat.copy()                     # a darker mark no sample lands on: page pixels tie at the maximum
    free = next((x, y) for y in range(30, 60) for x in range(40, 110)
                if all(abs(x - sx) > 6 or abs(y - sy) > 6 for sx, sy in sample_px))
    mark[free[1] - 3:free[1] + 3, free[0] - 3:free[0] + 3, :3] = (90, 5, 100)
    faint = flat.copy()                    # a brighter, faint one: ITS pixels are the maximum, tied up to rounding
    faint[..., :3] = (100, 100, 100)
    faint[free[1] - 3:free[1] + 3, free[0] - 3:free[0] + 3, :3] = (130, 186, 199)
    halves = flat.copy()                   # two tones, samples on both
    halves[:, w // 2:, :3] = (10, 10, 10)
    for name, page in (("flat", flat), ("mark", mark), ("faint", faint), ("halves", halves)):
        for slope, limit in ((10.0, 1000.0), (7.5, 1000.0)):       # integer and fractional limit/slope
            ref = oracle.ace_with_samples(page, xy, slope, limit, nb_threads=4)
            got = gpu.ace(page, slope, limit, n, seed=4)
            assert_exact(got, ref, "%s slope %g" % (name, slope))
    assert (gpu.ace(flat, 10, 1000, n, seed=4)[..., :3] == 0).all()
    assert (gpu.ace(mark, 10, 1000, n, seed=4)[0, 0, :3] == 255).all()


def test_randomised_sweep(gpu, oracle):
    """tools/fuzz_parity.py: random shapes, contents (noise, flat pages with
    spikes, blocks, ramps, binary textures, saturated pages) and parameters
    through all four calls."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import fuzz_parity
    bad = fuzz_parity.run(60, 1)
    assert not bad, bad
    # the same through the pinned-memory (band-pipelined) entry points
    bad = fuzz_parity.run(20, 2, pinned=True)
    assert not bad, bad


def test_ace_many_samples_chunked(gpu, oracle):
    # more samples than one shared-memory chunk holds (512)
    page = synthetic_page(96, 128, 6)
    xy = oracle.ace_draw_samples(700, 96, 128, 21)
    ref = oracle.ace_with_samples(page, xy, 10.0, 1000.0, nb_threads=8)
    got = gpu.ace(page, 10, 1000, 700, seed=21)
    assert_exact(got, ref, "ace 700 samples")
    # many more, on a page whose maximum is a near-tied set (a faint mark no sample may land on is
    # not guaranteed with this many samples: a two-tone page instead)
    page2 = np.full((64, 80, 4), 255, dtype=np.uint8)
    page2[..., :3] = 100
    page2[20:30, 30:50, :3] = (130, 186, 199)
    xy2 = oracle.ace_draw_samples(1300, 80, 64, 5)
    for slope, limit in ((10.0, 1000.0), (7.5, 1000.0)):
        ref2 = oracle.ace_with_samples(page2, xy2, slope, limit, nb_threads=8)
        got2 = gpu.ace(page2, slope, limit, 1300, seed=5)
        assert_exact(got2, ref2, "ace 1300 samples")


def test_ace_scores_within_their_bound(gpu, oracle):
    """The FP32 score planes of pass 1 against the oracle's FP64 scores: the
    error must stay inside the bound the later passes are built on (eps: what
    pf_ace.cu's ace_score_eps derives from the operation counts; restated here),
    tied scores must be exact, and pass 3's extremes must be the oracle's to the
    last bit."""
    import torch
    from libpillowfight_b200.distributed import CudaAceCompute, xy_tensor
    dev = torch.device("cuda:0")
    comp = CudaAceCompute()
    for w, h, n, idx in ((300, 200, 64, 8), (256, 512, 400, 3)):
        page = synthetic_page(w, h, idx)
        page[40:60, 50:90, :3] = (255, 255, 255)      # a saturated block: ties where every sample is darker by >= 100
        xy = oracle.ace_draw_samples(n, w, h, 5)
        ref = oracle.ace_scores(page, xy, 10.0, 1000.0, nb_threads=4)
        mm = comp.new_minmax(dev, w, h)
        tp, txy = torch.from_numpy(page).to(dev), xy_tensor(xy, dev)
        s = comp.scores(tp, txy, 10.0, 1000.0, (0, h), mm)
        comp.refine(tp, txy, 10.0, 1000.0, (0, h), s, mm)
        torch.cuda.synchronize()
        got = comp.score_planes(s, w, h).cpu().numpy().astype(np.float64)
        levels2 = (n + 15) // 16 + 2 * ((n + 511) // 512)
        eps = 4.0 * (15 + levels2) * 2.0 ** -24 * 1.01 + 3.8e-7 + 3.3e-7 + 1.0e-6
        err = float(np.nanmax(np.abs(got - ref)))
        assert err < eps, (err, eps)
        assert err < 0.25 * eps, (err, eps)            # (and in practice far inside it)
        ties = np.isin(got, (0.0, 1.0, -1.0))
        assert np.array_equal(got[ties], ref[ties])
        assert ties[:, 45:55, 55:85].any() or not np.isin(ref[:, 45:55, 55:85], (0.0, 1.0, -1.0)).any()
        mmh = mm.cpu().numpy()
        assert np.abs(mmh[:3] - np.nanmin(ref, axis=(1, 2))).max() < eps
        assert np.abs(mmh[3:6] - np.nanmax(ref, axis=(1, 2))).max() < eps
        assert np.array_equal(mmh[6:9], np.nanmin(ref, axis=(1, 2))), (mmh[6:9], np.nanmin(ref, axis=(1, 2)))
        assert np.array_equal(mmh[9:12], np.nanmax(ref, axis=(1, 2))), (mmh[9:12], np.nanmax(ref, axis=(1, 2)))


def test_ace_band_split_equals_whole(gpu, oracle):
    """Row bands computed separately (the multi-GPU split) give the same page."""
    import torch
    from libpillowfight_b200.distributed import CudaAceCompute, row_bands, xy_tensor
    w, h, n = 200, 151, 32
    page = synthetic_page(w, h, 9)
    xy = oracle.ace_draw_samples(n, w, h, 77)
    dev = torch.device("cuda:0")
    comp = CudaAceCompute()
    tp, txy = torch.from_numpy(page).to(dev), xy_tensor(xy, dev)
    mm = comp.new_minmax(dev, w, h)
    sc = comp.scores(tp, txy, 10.0, 1000.0, (0, h), mm)
    comp.refine(tp, txy, 10.0, 1000.0, (0, h), sc, mm)
    whole = comp.scale(tp, txy, 10.0, 1000.0, (0, h), sc, mm)
    mm2 = comp.new_minmax(dev, w, h)
    bands = row_bands(h, 3)
    parts = [comp.scores(tp, txy, 10.0, 1000.0, b, mm2) for b in bands]
    for b, p in zip(bands, parts):
        comp.refine(tp, txy, 10.0, 1000.0, b, p, mm2)
    split = torch.cat([comp.scale(tp, txy, 10.0, 1000.0, b, p, mm2) for b, p in zip(bands, parts)], dim=0)
    torch.cuda.synchronize()
    assert torch.equal(mm[:12], mm2[:12])
    assert torch.equal(whole, split)
    assert_exact(whole.cpu().numpy(), oracle.ace_with_samples(page, xy, 10.0, 1000.0, nb_threads=4), "whole")


def test_ace_bands_without_the_page(gpu, oracle):
    """pf_b200_dev_ace_scores_band / refine_band: a GPU that holds only its
    rows of the page and the samples' pixel values (what the multi-GPU split
    sends it) computes the same band as one that holds the page."""
    import torch
    from libpillowfight_b200.distributed import CudaAceCompute, row_bands, xy_tensor
    w, h, n = 200, 151, 32
    page = synthetic_page(w, h, 9)
    xy = oracle.ace_draw_samples(n, w, h, 77)
    dev = torch.device("cuda:0")
    comp = CudaAceCompute()
    tp, txy = torch.from_numpy(page).to(dev), xy_tensor(xy, dev)
    mm = comp.new_minmax(dev, w, h)
    sc = comp.scores(tp, txy, 10.0, 1000.0, (0, h), mm)
    comp.refine(tp, txy, 10.0, 1000.0, (0, h), sc, mm)
    whole = comp.scale(tp, txy, 10.0, 1000.0, (0, h), sc, mm)
    spx = comp.sample_pixels(tp, txy)
    assert np.array_equal(spx.cpu().numpy(), page[xy[1::2], xy[0::2]])
    mm2 = comp.new_minmax(dev, w, h)
    bands = row_bands(h, 3)
    # each band in its own buffer, surrounded by garbage that must not be read
    held = []
    for b0, b1 in bands:
        buf = torch.full(((b1 - b0 + 16) * w * 4,), 0xA5, dtype=torch.uint8, device=dev)
        rows = buf[8 * w * 4: (b1 - b0 + 8) * w * 4].view(b1 - b0, w, 4)
        rows.copy_(tp[b0:b1])
        held.append(rows)
    parts = [comp.scores_band(r, h, b, txy, spx, 10.0, 1000.0, mm2) for r, b in zip(held, bands)]
    for r, b, p in zip(held, bands, parts):
        comp.refine_band(r, h, b, txy, spx, 10.0, 1000.0, p, mm2)
    split = torch.cat([comp.scale_band(r, h, b, txy, spx, 10.0, 1000.0, p, mm2)
                       for r, b, p in zip(held, bands, parts)], dim=0)
    torch.cuda.synchronize()
    assert torch.equal(mm[:12], mm2[:12])
    assert torch.equal(whole, split)
    assert_exact(split.cpu().numpy(), oracle.ace_with_samples(page, xy, 10.0, 1000.0, nb_threads=4), "bands")
    # pf_b200_dev_ace_scale_to: every band stored into three whole-page images
    # at once (across GPUs these are peer buffers; here all on one device)
    from libpillowfight_b200 import _capi
    images = [torch.full((h, w, 4), 0x11, dtype=torch.uint8, device=dev) for _ in range(3)]
    st = torch.cuda.current_stream().cuda_stream
    # (the scores kernel leaves the fix-up lists empty; a block is rescaled once
    # per accumulation, so accumulate again)
    mm3 = comp.new_minmax(dev, w, h)
    parts = [comp.scores_band(r, h, b, txy, spx, 10.0, 1000.0, mm3) for r, b in zip(held, bands)]
    for r, b, p in zip(held, bands, parts):
        comp.refine_band(r, h, b, txy, spx, 10.0, 1000.0, p, mm3)
    for r, (b0, b1), p in zip(held, bands, parts):
        ptrs = (ctypes.c_void_p * 3)(*[im[b0:b1].data_ptr() for im in images])
        _capi.check(_capi.lib.pf_b200_dev_ace_scale_to(
            r.data_ptr(), w, h, b0, b1, txy.data_ptr(), spx.data_ptr(), n, 10.0, 1000.0,
            p.data_ptr(), mm3.data_ptr(), ptrs, 3, st))
    torch.cuda.synchronize()
    for im in images:
        assert torch.equal(im, whole)
    assert _capi.lib.pf_b200_dev_ace_scale_to(
        held[0].data_ptr(), w, h, bands[0][0], bands[0][1], txy.data_ptr(), spx.data_ptr(), n, 10.0, 1000.0,
        parts[0].data_ptr(), mm3.data_ptr(), ptrs, 17, st) == _capi.PF_B200_EINVAL


def test_empty_and_bad_arguments(gpu):
    """Zero-sized bitmaps, mismatched sizes, no samples, samples outside the
    page, tap counts out of range: an error, never a crash or a silent no-op."""
    from libpillowfight_b200 import _capi
    lib = _capi.lib
    buf = np.zeros((4, 4, 4), dtype=np.uint8)

    def bm(w, h, arr):
        b = _capi.PfBitmap()
        b.size.x, b.size.y, b.pixels = w, h, arr.ctypes.data
        return b

    for w, h in ((0, 4), (4, 0), (-1, 4)):
        b = bm(w, h, buf)
        assert lib.pf_b200_sobel(ctypes.byref(b), ctypes.byref(b)) == _capi.PF_B200_EINVAL
        assert lib.pf_b200_canny(ctypes.byref(b), ctypes.byref(b)) == _capi.PF_B200_EINVAL
        assert lib.pf_b200_gaussian(ctypes.byref(b), ctypes.byref(b), 2.0, 5) == _capi.PF_B200_EINVAL
    a, b = bm(4, 4, buf), bm(4, 3, buf)
    assert lib.pf_b200_sobel(ctypes.byref(a), ctypes.byref(b)) == _capi.PF_B200_EINVAL
    for n in (0, -3, 64, 1000):
        assert lib.pf_b200_gaussian(ctypes.byref(a), ctypes.byref(a), 2.0, n) == _capi.PF_B200_EINVAL
    xy = np.array([1, 1, 9, 1], dtype=np.int32)      # second sample outside a 4 x 4 page
    xyp = xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int))
    assert lib.pf_b200_ace(ctypes.byref(a), ctypes.byref(a), 2, xyp, 10.0, 1000.0) == _capi.PF_B200_EINVAL
    assert lib.pf_b200_ace(ctypes.byref(a), ctypes.byref(a), 0, xyp, 10.0, 1000.0) == _capi.PF_B200_EINVAL
    assert lib.pf_b200_ace(ctypes.byref(a), ctypes.byref(a), 1, xyp, -1.0, 1000.0) == _capi.PF_B200_EINVAL
    with pytest.raises(gpu.PillowfightB200Error):
        gpu.gaussian(buf, 2.0, 0)
    # the void entry points report instead of crashing and leave `out` alone
    out = np.full((4, 4, 4), 7, dtype=np.uint8)
    bz, bo = bm(0, 0, buf), bm(0, 0, out)
    lib.pf_apply_sobel(ctypes.byref(bz), ctypes.byref(bo))
    assert lib.pf_b200_last_error() == _capi.PF_B200_EINVAL and (out == 7).all()


def test_c_entry_points_and_aliases(gpu, oracle):
    """pf_apply_* (void, rand()-driven ACE) and the upstream spellings."""
    from libpillowfight_b200 import _capi
    page = synthetic_page(128, 96, 10)
    out = np.empty_like(page)
    bi = _capi.PfBitmap()
    bi.size.x, bi.size.y, bi.pixels = 128, 96, page.ctypes.data
    bo = _capi.PfBitmap()
    bo.size.x, bo.size.y, bo.pixels = 128, 96, out.ctypes.data
    lib = _capi.lib
    for name, ref in (("pf_apply_sobel", oracle.sobel(page)), ("pf_sobel", oracle.sobel(page)),
                      ("pf_apply_canny", oracle.canny(page)), ("pf_canny", oracle.canny(page))):
        out[:] = 0
        getattr(lib, name)(ctypes.byref(bi), ctypes.byref(bo))
        assert np.array_equal(out, ref), name
    for name in ("pf_apply_gaussian", "pf_gaussian"):
        out[:] = 0
        getattr(lib, name)(ctypes.byref(bi), ctypes.byref(bo), 2.0, 5)
        assert np.array_equal(out, oracle.gaussian(page, 2.0, 5)), name
    for name in ("pf_apply_ace", "pf_ace"):
        out[:] = 0
        oracle.srand(4242)
        getattr(lib, name)(ctypes.byref(bi), ctypes.byref(bo), 16, 10.0, 1000.0, 2)
        ref = oracle.ace(page, 10.0, 1000.0, 16, seed=4242, nb_threads=2)
        assert_exact(out, ref, name)


def test_in_place_and_pinned_buffers(gpu, oracle):
    from libpillowfight_b200 import _capi
    lib = _capi.lib
    page = synthetic_page(160, 100, 11)
    n = page.nbytes
    # in == out is allowed at the host-pointer boundary
    buf = page.copy()
    b = _capi.PfBitmap()
    b.size.x, b.size.y, b.pixels = 160, 100, buf.ctypes.data
    _capi.check(lib.pf_b200_sobel(ctypes.byref(b), ctypes.byref(b)))
    assert np.array_equal(buf, oracle.sobel(page))
    # pinned memory takes the direct-DMA path
    pin = lib.pf_b200_host_alloc(2 * n)
    assert pin
    try:
        ctypes.memmove(pin, page.ctypes.data, n)
        bi = _capi.PfBitmap()
        bi.size.x, bi.size.y, bi.pixels = 160, 100, pin
        bo = _capi.PfBitmap()
        bo.size.x, bo.size.y, bo.pixels = 160, 100, pin + n
        _capi.check(lib.pf_b200_canny(ctypes.byref(bi), ctypes.byref(bo)))
        got = np.ctypeslib.as_array((ctypes.c_uint8 * n).from_address(pin + n)).reshape(page.shape)
        assert np.array_equal(got, oracle.canny(page))
    finally:
        lib.pf_b200_host_free(pin)


def test_host_calls_side_by_side_on_several_threads(gpu, oracle):
    """Four host threads, one per call of a step, each making its call several
    times while the others run: a call that begins beside others moves its page
    in one DMA each way instead of row bands (pf_capi.cu, CallScope).  Every
    result of every round is compared with the oracle; Sobel also in place."""
    import threading
    from libpillowfight_b200 import _capi
    lib = _capi.lib
    w, h = 700, 1500
    page = synthetic_page(w, h, 21)
    n = page.nbytes
    pin = lib.pf_b200_host_alloc(6 * n)
    assert pin
    try:
        def view(k):
            return np.ctypeslib.as_array((ctypes.c_uint8 * n).from_address(pin + k * n)).reshape(page.shape)

        def bm(k):
            b = _capi.PfBitmap()
            b.size.x, b.size.y, b.pixels = w, h, pin + k * n
            return b

        view(0)[:] = page
        xy = oracle.ace_draw_samples(40, w, h, 5)
        xyp = xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int))
        refs = {"ace": oracle.ace_with_samples(page, xy, 10.0, 1000.0, nb_threads=8),
                "gaussian": oracle.gaussian(page, 2.0, 5), "sobel": oracle.sobel(page),
                "canny": oracle.canny(page)}
        b_in = bm(0)
        outs = {"ace": 1, "gaussian": 2, "sobel": 3, "canny": 4}
        calls = {
            "ace": lambda: lib.pf_b200_ace(ctypes.byref(b_in), ctypes.byref(bm(1)), 40, xyp, 10.0, 1000.0),
            "gaussian": lambda: lib.pf_b200_gaussian(ctypes.byref(b_in), ctypes.byref(bm(2)), 2.0, 5),
            "canny": lambda: lib.pf_b200_canny(ctypes.byref(b_in), ctypes.byref(bm(4))),
        }
        bad = []
        start = threading.Barrier(4)

        def worker(name):
            try:
                start.wait()
                for rnd in range(6):
                    if name == "sobel":             # in place on a private copy of the page
                        view(3)[:] = page
                        _capi.check(lib.pf_b200_sobel(ctypes.byref(bm(3)), ctypes.byref(bm(3))))
                    else:
                        view(outs[name])[:] = 0x5A
                        _capi.check(calls[name]())
                    if not np.array_equal(view(outs[name]), refs[name]):
                        bad.append("%s, round %d" % (name, rnd))
            except Exception as exc:            # noqa: BLE001
                bad.append("%s: %r" % (name, exc))

        threads = [threading.Thread(target=worker, args=(name,)) for name in outs]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        assert not bad, bad
    finally:
        lib.pf_b200_host_free(pin)


def test_streamed_paths_on_pinned_pages(gpu, oracle):
    """Pinned caller memory and >= 512 rows: the page is uploaded, processed
    and returned band by band on three streams.  Same results, also in place."""
    from libpillowfight_b200 import _capi
    lib = _capi.lib
    w, h = 700, 1500                     # 6 bands of 256 rows, the last one short
    page = synthetic_page(w, h, 13)
    n = page.nbytes
    pin = lib.pf_b200_host_alloc(2 * n)
    assert pin
    try:
        def view(k):
            return np.ctypeslib.as_array((ctypes.c_uint8 * n).from_address(pin + k * n)).reshape(page.shape)

        def bm(k):
            b = _capi.PfBitmap()
            b.size.x, b.size.y, b.pixels = w, h, pin + k * n
            return b

        src, dst = view(0), view(1)
        b_src, b_dst = bm(0), bm(1)
        src[:] = page
        _capi.check(lib.pf_b200_gaussian(ctypes.byref(b_src), ctypes.byref(b_dst), 2.0, 5))
        ref_g = oracle.gaussian(page, 2.0, 5)
        assert np.array_equal(dst, ref_g)
        _capi.check(lib.pf_b200_gaussian(ctypes.byref(b_src), ctypes.byref(b_dst), 3.0, 9))
        assert np.array_equal(dst, oracle.gaussian(page, 3.0, 9))
        _capi.check(lib.pf_b200_sobel(ctypes.byref(b_src), ctypes.byref(b_dst)))
        assert np.array_equal(dst, oracle.sobel(page))
        # Canny: bands are classified and returned as they arrive, the weak
        # pixels are written into the caller's page again after hysteresis
        ref_c = oracle.canny(page)
        dst[:] = 0x5A
        _capi.check(lib.pf_b200_canny(ctypes.byref(b_src), ctypes.byref(b_dst)))
        assert np.array_equal(dst, ref_c)
        # ... and a page where most pixels are weak or strong (second full download)
        noisy = uniform_random_page(w, h, seed=3)
        src[:] = noisy
        _capi.check(lib.pf_b200_canny(ctypes.byref(b_src), ctypes.byref(b_dst)))
        assert np.array_equal(dst, oracle.canny(noisy))
        # ... and long weak chains that only a far-away strong pixel keeps
        chain = np.zeros((h, w, 4), dtype=np.uint8)
        chain[..., 3] = 255
        chain[:, 300:, :3] = 44                 # a weak vertical edge down the whole page
        chain[700:703, 290:310, :3] = 255       # one strong spot touching it
        src[:] = chain
        _capi.check(lib.pf_b200_canny(ctypes.byref(b_src), ctypes.byref(b_src)))       # in place
        assert np.array_equal(src, oracle.canny(chain))
        assert (src[..., 0] == 255).sum() > h   # the edge was kept
        src[:] = page
        xy = oracle.ace_draw_samples(60, w, h, 31)
        _capi.check(lib.pf_b200_ace(ctypes.byref(b_src), ctypes.byref(b_dst), 60,
                                    xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), 10.0, 1000.0))
        ref_a = oracle.ace_with_samples(page, xy, 10.0, 1000.0, nb_threads=8)
        assert_exact(dst, ref_a, "streamed ace")
        # the streamed and the plain path run the same kernels on the same tiles
        assert np.array_equal(dst, gpu.ace(page, 10, 1000, 60, seed=31))
        # in place
        _capi.check(lib.pf_b200_sobel(ctypes.byref(b_src), ctypes.byref(b_src)))
        assert np.array_equal(src, oracle.sobel(page))
        src[:] = page
        _capi.check(lib.pf_b200_gaussian(ctypes.byref(b_src), ctypes.byref(b_src), 2.0, 5))
        assert np.array_equal(src, ref_g)
        src[:] = page
        _capi.check(lib.pf_b200_ace(ctypes.byref(b_src), ctypes.byref(b_src), 60,
                                    xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), 10.0, 1000.0))
        assert np.array_equal(src, dst)        # same as the out-of-place ACE above
    finally:
        lib.pf_b200_host_free(pin)


def test_pageable_band_pipeline(gpu, oracle):
    """Ordinary (malloc) memory on both sides, a page of 8 MB or more: host
    threads stage bands in and out around the band pipeline.  Also in place,
    and with a height that leaves a short last band."""
    from libpillowfight_b200 import _capi
    lib = _capi.lib
    w, h = 1704, 1237                    # 8.4 MB, 16 bands of 78 -> 128 rows: 10 bands, the last short
    page = synthetic_page(w, h, 21)

    def bm(a):
        b = _capi.PfBitmap()
        b.size.x, b.size.y, b.pixels = w, h, a.ctypes.data
        return b

    out = np.full_like(page, 0x5A)
    b_in, b_out = bm(page), bm(out)
    ref_g, ref_s = oracle.gaussian(page, 2.0, 5), oracle.sobel(page)
    for _ in range(3):                   # repeated: staging areas are reused
        _capi.check(lib.pf_b200_gaussian(ctypes.byref(b_in), ctypes.byref(b_out), 2.0, 5))
        assert np.array_equal(out, ref_g)
        _capi.check(lib.pf_b200_sobel(ctypes.byref(b_in), ctypes.byref(b_out)))
        assert np.array_equal(out, ref_s)
    _capi.check(lib.pf_b200_gaussian(ctypes.byref(b_in), ctypes.byref(b_out), 3.0, 9))
    assert np.array_equal(out, oracle.gaussian(page, 3.0, 9))
    work = page.copy()
    b_work = bm(work)
    _capi.check(lib.pf_b200_sobel(ctypes.byref(b_work), ctypes.byref(b_work)))        # in place
    assert np.array_equal(work, ref_s)
    work[:] = page
    _capi.check(lib.pf_b200_gaussian(ctypes.byref(b_work), ctypes.byref(b_work), 2.0, 5))
    assert np.array_equal(work, ref_g)
    # Canny on the same memory (plain staging)
    _capi.check(lib.pf_b200_canny(ctypes.byref(b_in), ctypes.byref(b_out)))
    assert np.array_equal(out, oracle.canny(page))
    # ACE: bands accumulated as they are staged in, rescaled bands staged out
    xy = oracle.ace_draw_samples(40, w, h, 17)
    xyp = xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int))
    ref_a = oracle.ace_with_samples(page, xy, 10.0, 1000.0, nb_threads=8)
    for _ in range(2):
        out[:] = 0x5A
        _capi.check(lib.pf_b200_ace(ctypes.byref(b_in), ctypes.byref(b_out), 40, xyp, 10.0, 1000.0))
        assert_exact(out, ref_a, "pageable ace")
    assert np.array_equal(out, gpu.ace(page, 10, 1000, 40, seed=17))     # same kernels, same tiles
    work[:] = page
    _capi.check(lib.pf_b200_ace(ctypes.byref(b_work), ctypes.byref(b_work), 40, xyp, 10.0, 1000.0))
    assert np.array_equal(work, out)


def test_device_batch_entry_points(gpu, oracle):
    import torch
    from libpillowfight_b200.distributed import CudaChain
    dev = torch.device("cuda:0")
    pages = np.stack([synthetic_page(132, 90, i) for i in range(5)])
    tin = torch.from_numpy(pages).to(dev)
    chain = CudaChain()
    g = chain.gaussian(tin, torch.empty_like(tin))
    s = chain.sobel(g, torch.empty_like(tin))
    c1 = chain.canny(tin, torch.empty_like(tin), pages_per_launch=1)
    c5 = chain.canny(tin, torch.empty_like(tin), pages_per_launch=5)
    torch.cuda.synchronize()
    for i in range(5):
        rg = oracle.gaussian(pages[i], 2.0, 5)
        assert np.array_equal(g[i].cpu().numpy(), rg)
        assert np.array_equal(s[i].cpu().numpy(), oracle.sobel(rg))
        rc = oracle.canny(pages[i])
        assert np.array_equal(c1[i].cpu().numpy(), rc)
        assert np.array_equal(c5[i].cpu().numpy(), rc)


def test_device_ace_whole_page(gpu, oracle):
    """pf_b200_dev_ace: the ACE passes for a page resident in HBM, one call, one
    caller-owned work area; same result as the host-pointer entry point."""
    import torch
    from libpillowfight_b200 import _capi
    from libpillowfight_b200.distributed import xy_tensor
    lib = _capi.lib
    dev = torch.device("cuda:0")
    w, h, n = 212, 160, 50
    page = synthetic_page(w, h, 12)
    xy = oracle.ace_draw_samples(n, w, h, 9)
    tin, txy = torch.from_numpy(page).to(dev), xy_tensor(xy, dev)
    out = torch.empty_like(tin)
    need = lib.pf_b200_ace_workspace_bytes(w, h)
    assert need >= 16 * w * h
    work = torch.empty(need, dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    _capi.check(lib.pf_b200_dev_ace(tin.data_ptr(), out.data_ptr(), w, h, txy.data_ptr(), n, 10.0, 1000.0,
                                    work.data_ptr(), need, st))
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    assert_exact(got, oracle.ace_with_samples(page, xy, 10.0, 1000.0, nb_threads=4), "dev ace")
    assert np.array_equal(got, gpu.ace(page, 10, 1000, n, seed=9))
    # too small a work area, in place: refused
    assert lib.pf_b200_dev_ace(tin.data_ptr(), out.data_ptr(), w, h, txy.data_ptr(), n, 10.0, 1000.0,
                               work.data_ptr(), need - 1, st) == _capi.PF_B200_EINVAL
    assert lib.pf_b200_dev_ace(tin.data_ptr(), tin.data_ptr(), w, h, txy.data_ptr(), n, 10.0, 1000.0,
                               work.data_ptr(), need, st) == _capi.PF_B200_EINVAL


@pytest.mark.parametrize("w,h", [(5, 4), (33, 17), (124, 61), (128, 40), (257, 129), (620, 300)])
def test_no_writes_outside_the_output_page(gpu, oracle, w, h):
    """compute-sanitizer is not available on this pool, so: guard bands of a
    known pattern in front of and behind every output buffer (and the Canny
    scratch) must come back untouched, and results must still match."""
    import torch
    from libpillowfight_b200 import _capi
    lib, chk = _capi.lib, _capi.check
    dev = torch.device("cuda:0")
    page = uniform_random_page(w, h, seed=3 * w + h)
    guard = 4096                                   # bytes, keeps 256-byte alignment
    n = w * h * 4

    def guarded(nbytes):
        buf = torch.full((guard + nbytes + guard,), 0xA5, dtype=torch.uint8, device=dev)
        return buf, buf[guard:guard + nbytes]

    def intact(buf, nbytes):
        return bool((buf[:guard] == 0xA5).all()) and bool((buf[guard + nbytes:] == 0xA5).all())

    tin = torch.from_numpy(page).to(dev)
    st = torch.cuda.current_stream().cuda_stream
    for name in ("gaussian", "sobel", "canny", "ace"):
        obuf, out = guarded(n)
        if name == "gaussian":
            chk(lib.pf_b200_dev_gaussian(tin.data_ptr(), out.data_ptr(), w, h, 2.0, 5, st))
            ref = oracle.gaussian(page, 2.0, 5)
        elif name == "sobel":
            chk(lib.pf_b200_dev_sobel(tin.data_ptr(), out.data_ptr(), w, h, st))
            ref = oracle.sobel(page)
        elif name == "canny":
            sbytes = lib.pf_b200_canny_scratch_bytes(w, h)
            sbuf, scratch = guarded(sbytes)
            chk(lib.pf_b200_dev_canny(tin.data_ptr(), out.data_ptr(), w, h, 1,
                                      scratch.data_ptr(), sbytes, st))
            ref = oracle.canny(page)
        else:
            if h < 5:
                continue
            xy = oracle.ace_draw_samples(10, w, h, 5)
            from libpillowfight_b200.distributed import xy_tensor
            txy = xy_tensor(xy, dev)
            sc_bytes = lib.pf_b200_ace_scores_bytes(w, h)
            sc_buf, sc = guarded(sc_bytes)
            mm_bytes = lib.pf_b200_ace_minmax_bytes(w, h)
            mm_buf, mm = guarded(mm_bytes)
            chk(lib.pf_b200_dev_ace_minmax_init(mm.data_ptr(), w, h, st))
            chk(lib.pf_b200_dev_ace_scores(tin.data_ptr(), w, h, 0, h, txy.data_ptr(), 10,
                                           10.0, 1000.0, sc.data_ptr(), mm.data_ptr(), st))
            chk(lib.pf_b200_dev_ace_refine(tin.data_ptr(), w, h, 0, h, txy.data_ptr(), 10,
                                           10.0, 1000.0, sc.data_ptr(), mm.data_ptr(), st))
            chk(lib.pf_b200_dev_ace_scale(tin.data_ptr(), w, h, 0, h, txy.data_ptr(), 10,
                                          10.0, 1000.0, sc.data_ptr(), mm.data_ptr(), out.data_ptr(), st))
            ref = oracle.ace_with_samples(page, xy, 10.0, 1000.0)
        torch.cuda.synchronize()
        assert intact(obuf, n), "%s wrote outside its output page" % name
        if name == "canny":
            assert intact(sbuf, sbytes), "canny wrote outside its scratch"
        if name == "ace":
            assert intact(sc_buf, sc_bytes), "ace wrote outside its score block"
            assert intact(mm_buf, mm_bytes), "ace wrote outside its min/max block"
            assert_exact(out.cpu().numpy().reshape(h, w, 4), ref, "ace guarded")
        else:
            assert np.array_equal(out.cpu().numpy().reshape(h, w, 4), ref), name


def test_concurrent_host_threads(gpu, oracle):
    """Each host thread has its own context (stream, buffers): four threads
    calling different entry points at once get the right answers."""
    import threading
    page = synthetic_page(400, 300, 14)
    xy_seed = 9
    refs = {"gaussian": oracle.gaussian(page, 2.0, 5), "sobel": oracle.sobel(page),
            "canny": oracle.canny(page)}
    errors = []

    def work(name):
        try:
            for _ in range(5):
                got = getattr(gpu, name)(page)
                if not np.array_equal(got, refs[name]):
                    errors.append(name)
        except Exception as e:   # noqa: BLE001
            errors.append("%s: %r" % (name, e))

    def work_ace():
        try:
            a = gpu.ace(page, 10, 1000, 20, seed=xy_seed)
            b = gpu.ace(page, 10, 1000, 20, seed=xy_seed)
            if not np.array_equal(a, b):
                errors.append("ace not reproducible")
        except Exception as e:   # noqa: BLE001
            errors.append("ace: %r" % (e,))

    threads = [threading.Thread(target=work, args=(n,)) for n in refs] + [threading.Thread(target=work_ace)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors


def test_pil_images_round_trip(gpu, oracle):
    from PIL import Image
    page = synthetic_page(90, 70, 12)
    img = Image.fromarray(page[..., :3], "RGB")          # converted to RGBA like upstream
    out = gpu.sobel(img)
    assert out.mode == "RGBA" and out.size == (90, 70)
    assert np.array_equal(np.asarray(out), oracle.sobel(page))
    # an RGBA image goes in as it is; the result image wraps the (pooled,
    # pinned) result buffer: it must outlive other calls and survive edits
    import gc
    rgba = Image.fromarray(page, "RGBA")
    outs = [gpu.gaussian(rgba), gpu.canny(rgba), gpu.ace(rgba, samples=20, seed=3)]
    for _ in range(6):
        gpu.sobel(rgba)                                   # churn the buffer pool
    gc.collect()
    assert np.array_equal(np.asarray(outs[0]), oracle.gaussian(page, 2.0, 5))
    assert np.array_equal(np.asarray(outs[1]), oracle.canny(page))
    assert_exact(np.asarray(outs[2]), oracle.ace_with_samples(
        page, oracle.ace_draw_samples(20, 90, 70, 3), 10.0, 1000.0), "ace PIL")
    edited = outs[1].copy()
    edited.putpixel((0, 0), (1, 2, 3, 4))
    outs[1].putpixel((1, 1), (5, 6, 7, 8))               # Pillow copies before writing
    assert outs[1].getpixel((1, 1)) == (5, 6, 7, 8)
    assert np.array_equal(np.asarray(gpu.canny(rgba)), oracle.canny(page))
    # other modes are converted to RGBA first, as upstream does
    gray = rgba.convert("L")
    as_rgba = np.asarray(gray.convert("RGBA"))
    assert np.array_equal(np.asarray(gpu.sobel(gray)), oracle.sobel(as_rgba))
    # a tall image (the pinned, band-pipelined path) from an RGB image
    tall = synthetic_page(260, 700, 5)
    assert np.array_equal(np.asarray(gpu.gaussian(Image.fromarray(tall[..., :3], "RGB"))),
                          oracle.gaussian(tall, 2.0, 5))


# ---- BASELINE.json's full sizes ----------------------------------------

def test_full_a4_chain_against_oracle(gpu, oracle):
    """configs[1] and [2]: Gaussian then Sobel, and Canny, on the A4 300-dpi page."""
    page = synthetic_page(*A4_300DPI, page_index=0)
    g = gpu.gaussian(page, 2.0, 5)
    rg = oracle.gaussian(page, 2.0, 5)
    assert np.array_equal(g, rg)
    assert np.array_equal(gpu.sobel(g), oracle.sobel(rg))
    assert np.array_equal(gpu.canny(page), oracle.canny(page))


def test_full_a4_ace_against_oracle(gpu, oracle):
    """configs[0]: ACE, A4 300 dpi, slope 10, limit 1000, 100 samples, fixed seed."""
    import os
    page = synthetic_page(*A4_300DPI, page_index=0)
    w, h = A4_300DPI
    xy = oracle.ace_draw_samples(100, w, h, 12345)
    ref = oracle.ace_with_samples(page, xy, 10.0, 1000.0, nb_threads=os.cpu_count() or 8)
    got = gpu.ace(page, 10, 1000, 100, seed=12345)
    assert_exact(got, ref, "ace A4")


def test_size_independent_properties(gpu):
    """Properties that hold at any size, checked at A4 without the oracle."""
    page = synthetic_page(*A4_300DPI, page_index=1)
    c = gpu.canny(page)
    assert set(np.unique(c[..., 0])) <= {0, 255}
    assert (c[..., 0] == c[..., 1]).all() and (c[..., 1] == c[..., 2]).all()
    assert (c[..., 3] == 255).all()
    # determinism: hysteresis is a fixed point, independent of scheduling
    assert np.array_equal(c, gpu.canny(page))
    # Sobel of a constant page is zero away from the border
    flat = np.full((500, 700, 4), 77, dtype=np.uint8)
    s = gpu.sobel(flat)
    assert not s[1:-1, 1:-1, :3].any() and s[0, 5, 0] > 0
    # ACE output spans the full range per channel and is reproducible per seed
    # (the top value is trunc(fl(255*D)/D) with D = max-min: 255, or 254 when
    # the product rounds down -- the reference formula has the same property)
    a = gpu.ace(page, 10, 1000, 50, seed=3)
    for ch in range(3):
        assert a[..., ch].min() == 0 and a[..., ch].max() >= 254
    assert np.array_equal(a, gpu.ace(page, 10, 1000, 50, seed=3))
    assert not np.array_equal(a, gpu.ace(page, 10, 1000, 50, seed=4))


def test_config4_ace_600dpi_400_samples(gpu, oracle):
    """BASELINE.json config 4 at full size: ACE on 4960 x 7016 with 400 samples,
    through pf_b200_ace (host pointers) and through the row-band entry points the
    multi-GPU split uses (here: three bands on one GPU), against the oracle."""
    import os
    import torch
    from libpillowfight_b200.distributed import CudaAceCompute, ace_sharded, row_bands, xy_tensor
    from libpillowfight_b200.pages import A4_600DPI
    w, h = A4_600DPI
    small = synthetic_page(*A4_300DPI, page_index=0)
    page = np.ascontiguousarray(np.repeat(np.repeat(small, 2, axis=0), 2, axis=1)[:h, :w])
    xy = oracle.ace_draw_samples(400, w, h, 12345)
    ref = oracle.ace_with_samples(page, xy, 10.0, 1000.0, nb_threads=os.cpu_count() or 8)
    assert_exact(gpu.ace(page, 10, 1000, 400, seed=12345), ref, "ace 600 dpi, host pointers")
    dev = torch.device("cuda:0")
    comp = CudaAceCompute()
    tp, txy = torch.from_numpy(page).to(dev), xy_tensor(xy, dev)
    assert_exact(ace_sharded(tp, txy, 10.0, 1000.0, comp).cpu().numpy(), ref, "ace 600 dpi, ace_sharded")
    mm = comp.new_minmax(dev, w, h)
    spx = comp.sample_pixels(tp, txy)
    bands = row_bands(h, 3)
    held = [tp[b0:b1].clone() for b0, b1 in bands]
    parts = [comp.scores_band(r, h, b, txy, spx, 10.0, 1000.0, mm) for r, b in zip(held, bands)]
    for r, b, p in zip(held, bands, parts):
        comp.refine_band(r, h, b, txy, spx, 10.0, 1000.0, p, mm)
    split = torch.cat([comp.scale_band(r, h, b, txy, spx, 10.0, 1000.0, p, mm)
                       for r, b, p in zip(held, bands, parts)], dim=0)
    assert_exact(split.cpu().numpy(), ref, "ace 600 dpi, three row bands")


def test_config5_page_batch_at_a4(gpu, oracle):
    """BASELINE.json config 5 at one GPU's share: 64 A4 pages through the
    batched Gaussian -> Sobel and Canny entry points; eight of the pages
    (first, last and some in between) are compared with the oracle bit for bit."""
    import torch
    from concurrent.futures import ThreadPoolExecutor
    from libpillowfight_b200.distributed import CudaChain
    dev = torch.device("cuda:0")
    w, h = A4_300DPI
    nb = 64
    base = [synthetic_page(w, h, i) for i in range(4)]
    batch = torch.empty((nb, h, w, 4), dtype=torch.uint8, device=dev)
    for i in range(nb):
        batch[i] = torch.from_numpy(np.roll(base[i % 4], 53 * i, axis=0)).to(dev)
    chain = CudaChain()
    g = chain.gaussian(batch, torch.empty_like(batch))
    s = chain.sobel(g, torch.empty_like(batch))
    c = chain.canny(batch, torch.empty_like(batch), pages_per_launch=8)
    torch.cuda.synchronize()
    picks = [0, 1, 7, 8, 31, 40, 62, 63]

    def check(i):
        page = np.roll(base[i % 4], 53 * i, axis=0)
        rg = oracle.gaussian(page, 2.0, 5)
        return (i, np.array_equal(g[i].cpu().numpy(), rg),
                np.array_equal(s[i].cpu().numpy(), oracle.sobel(rg)),
                np.array_equal(c[i].cpu().numpy(), oracle.canny(page)))

    with ThreadPoolExecutor(max_workers=8) as ex:
        for i, ok_g, ok_s, ok_c in ex.map(check, picks):
            assert ok_g and ok_s and ok_c, "page %d: gaussian %s sobel %s canny %s" % (i, ok_g, ok_s, ok_c)


def test_page_step_one_upload(gpu, oracle):
    """pf_b200_page_step: ACE, Gaussian, Sobel of the blurred page and Canny of a
    page from one upload -- on pinned pages (the overlapped path), with subsets
    of the outputs, and on ordinary memory (served by the drop-in calls)."""
    from libpillowfight_b200 import _capi
    lib = _capi.lib
    w, h, n = 620, 877, 60
    page = synthetic_page(w, h, 3)
    xy = oracle.ace_draw_samples(n, w, h, 31)
    xyp = xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int))
    rg = oracle.gaussian(page, 2.0, 5)
    refs = [oracle.ace_with_samples(page, xy, 10.0, 1000.0, nb_threads=8), rg, oracle.sobel(rg),
            oracle.canny(page)]
    nbytes = page.nbytes
    pin = lib.pf_b200_host_alloc(5 * nbytes)
    assert pin
    try:
        views = [np.ctypeslib.as_array((ctypes.c_uint8 * nbytes).from_address(pin + k * nbytes)).reshape(h, w, 4)
                 for k in range(5)]
        views[0][:] = page

        def bm(ptr):
            b = _capi.PfBitmap()
            b.size.x, b.size.y, b.pixels = w, h, ptr
            return b

        bms = [bm(pin + k * nbytes) for k in range(5)]
        for wanted in ((1, 1, 1, 1), (0, 0, 1, 0), (1, 0, 0, 1), (0, 1, 0, 0)):
            for v in views[1:]:
                v[:] = 0x5A
            outs = [ctypes.byref(bms[k + 1]) if wanted[k] else None for k in range(4)]
            _capi.check(lib.pf_b200_page_step(ctypes.byref(bms[0]), *outs, n, xyp, 10.0, 1000.0, 2.0, 5))
            for k in range(4):
                if wanted[k]:
                    assert_exact(views[k + 1], refs[k], "page step %s output %d" % (wanted, k))
                else:
                    assert (views[k + 1] == 0x5A).all()
        # refused: an output that is the input, two outputs on one page, nothing asked for
        assert lib.pf_b200_page_step(ctypes.byref(bms[0]), ctypes.byref(bms[0]), None, None, None,
                                     n, xyp, 10.0, 1000.0, 2.0, 5) == _capi.PF_B200_EINVAL
        assert lib.pf_b200_page_step(ctypes.byref(bms[0]), None, ctypes.byref(bms[1]), ctypes.byref(bms[1]), None,
                                     n, xyp, 10.0, 1000.0, 2.0, 5) == _capi.PF_B200_EINVAL
        assert lib.pf_b200_page_step(ctypes.byref(bms[0]), None, None, None, None,
                                     n, xyp, 10.0, 1000.0, 2.0, 5) == _capi.PF_B200_EINVAL
    finally:
        lib.pf_b200_host_free(pin)
    # ordinary memory and a small page: the same answers through the drop-in calls
    small = synthetic_page(150, 90, 4)
    xs = oracle.ace_draw_samples(20, 150, 90, 5)
    outs = [np.empty_like(small) for _ in range(4)]

    def bma(a):
        b = _capi.PfBitmap()
        b.size.x, b.size.y, b.pixels = a.shape[1], a.shape[0], a.ctypes.data
        return b

    b_in, b_outs = bma(small), [bma(o) for o in outs]
    _capi.check(lib.pf_b200_page_step(ctypes.byref(b_in), ctypes.byref(b_outs[0]), None, ctypes.byref(b_outs[2]),
                                      ctypes.byref(b_outs[3]), 20,
                                      xs.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), 10.0, 1000.0, 2.0, 5))
    sg = oracle.gaussian(small, 2.0, 5)
    assert_exact(outs[0], oracle.ace_with_samples(small, xs, 10.0, 1000.0), "page step small ace")
    assert np.array_equal(outs[2], oracle.sobel(sg))
    assert np.array_equal(outs[3], oracle.canny(small))


def test_ace_multi_gpu_from_one_process(gpu, oracle):
    """pf_b200_ace_multi_gpu: row bands on every visible GPU of this process
    (one GPU: the call serves the page through pf_b200_ace), against the oracle."""
    from libpillowfight_b200 import _capi
    lib = _capi.lib
    w, h, n = 620, 877, 60
    page = synthetic_page(w, h, 6)
    xy = oracle.ace_draw_samples(n, w, h, 9)
    xyp = xy.ctypes.data_as(ctypes.POINTER(ctypes.c_int))
    ref = oracle.ace_with_samples(page, xy, 10.0, 1000.0, nb_threads=8)
    nbytes = page.nbytes
    pin = lib.pf_b200_host_alloc(2 * nbytes)
    assert pin
    try:
        src = np.ctypeslib.as_array((ctypes.c_uint8 * nbytes).from_address(pin)).reshape(h, w, 4)
        dst = np.ctypeslib.as_array((ctypes.c_uint8 * nbytes).from_address(pin + nbytes)).reshape(h, w, 4)
        src[:] = page
        bi, bo = _capi.PfBitmap(), _capi.PfBitmap()
        bi.size.x, bi.size.y, bi.pixels = w, h, pin
        bo.size.x, bo.size.y, bo.pixels = w, h, pin + nbytes
        for nd in (0, 1, 2, 3):
            dst[:] = 0x5A
            _capi.check(lib.pf_b200_ace_multi_gpu(ctypes.byref(bi), ctypes.byref(bo), n, xyp, 10.0, 1000.0, None, nd))
            assert_exact(dst, ref, "ace on %d devices" % nd)
        bad = (ctypes.c_int * 2)(0, 0)
        if gpu.device_count() >= 2:
            assert lib.pf_b200_ace_multi_gpu(ctypes.byref(bi), ctypes.byref(bo), n, xyp, 10.0, 1000.0, bad, 2) \
                == _capi.PF_B200_EINVAL
        bad[0] = 99
        assert lib.pf_b200_ace_multi_gpu(ctypes.byref(bi), ctypes.byref(bo), n, xyp, 10.0, 1000.0, bad, 1) \
            == _capi.PF_B200_EINVAL
    finally:
        lib.pf_b200_host_free(pin)


def test_ace_low_contrast_pages(gpu, oracle):
    """Pages whose scores span a tiny range: the rescale magnifies the fast
    pass's error until every pixel is 'uncertain' (the guard band reaches half
    an output step) and the whole page goes through the exact passes; pages
    that are flat except for a few pixels; a smooth ramp (neighbouring pixels'
    scores closer than the error bound).  All must still equal the oracle."""
    rng = np.random.default_rng(11)
    w, h, n = 300, 200, 24
    xy = oracle.ace_draw_samples(n, w, h, 6)
    pages = {}
    blank = np.full((h, w, 4), 255, dtype=np.uint8)
    blank[..., :3] = 200 + rng.integers(0, 2, size=(h, w, 3), dtype=np.uint8)
    pages["one level of noise"] = blank
    faint = np.full((h, w, 4), 255, dtype=np.uint8)
    faint[..., :3] = (90, 120, 150)
    faint[50, 60, :3] = (91, 120, 150)
    faint[150, 200, :3] = (90, 119, 150)
    pages["flat with two pixels off by one"] = faint
    ramp = np.full((h, w, 4), 255, dtype=np.uint8)
    ramp[..., :3] = (np.arange(w)[None, :, None] * 40 // w + 100).astype(np.uint8)
    pages["smooth ramp"] = ramp
    dark = np.zeros((h, w, 4), dtype=np.uint8)
    dark[..., 3] = 255
    dark[..., :3] = rng.integers(0, 3, size=(h, w, 3), dtype=np.uint8)
    pages["near black"] = dark
    for name, page in pages.items():
        for slope, limit in ((10.0, 1000.0), (3.3, 841.5)):
            ref = oracle.ace_with_samples(page, xy, slope, limit, nb_threads=8)
            assert_exact(gpu.ace(page, slope, limit, n, seed=6), ref, "%s, slope %g" % (name, slope))


def test_ace_error_constants_hold_on_this_device(gpu):
    """The two hardware constants ACE's error bounds assume (pf_ace.cu:
    ace_score_eps, ace_accurate_eps), measured exhaustively on the device the
    tests run on."""
    from libpillowfight_b200 import _capi
    e32 = _capi.lib.pf_b200_selftest_rsqrt_error(0)
    e64 = _capi.lib.pf_b200_selftest_rsqrt_error(1)
    assert 0.0 < e32 <= 1.25e-7, e32
    assert 0.0 < e64 <= 2.0 ** -48, e64
    print("rsqrt errors: FP32 %.3e (2^%.2f), FP64 after two Newton steps %.3e (2^%.2f)" % (
        e32, np.log2(e32), e64, np.log2(e64)))
