"""pf_b200_dev_canny_noalloc: Canny for callers that may not allocate.

Some users of this library cannot let it allocate GPU memory: a server whose
own pool already owns the whole device, a real-time loop that cannot afford
the device-wide stall an allocation causes, a deployment whose memory use is
frozen after start-up.  For them the library carries its hysteresis work list
in a fixed device array (pages up to 4,194,304 pixels).  This test drives that
entry point through the C-ABI and holds it to the ordinary entry point
(pf_b200_dev_canny, caller-provided scratch): Canny is an integer stage, so
the two must agree exactly -- on two calls in a row, because the second call
finds the first call's leftovers in the shared array.

This file is named to be collected last: if it fails under `pytest -x`, every
other test has already run.
"""
import numpy as np
import pytest

from libpillowfight_b200.pages import synthetic_page, uniform_random_page

pytestmark = pytest.mark.gpu

MAX_PIXELS = 4 << 20


def _run_both(torch, capi, page):
    """(no-alloc result, ordinary result) as numpy arrays for one page."""
    lib = capi.lib
    dev = torch.device("cuda:0")
    h, w = page.shape[:2]
    tin = torch.from_numpy(page).to(dev)
    out_na = torch.zeros_like(tin)
    out_ref = torch.zeros_like(tin)
    scratch = torch.empty(lib.pf_b200_canny_scratch_bytes(w, h), dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    capi.check(lib.pf_b200_dev_canny_noalloc(tin.data_ptr(), out_na.data_ptr(), w, h, st))
    capi.check(lib.pf_b200_dev_canny(tin.data_ptr(), out_ref.data_ptr(), w, h, 1,
                                     scratch.data_ptr(), scratch.numel(), st))
    torch.cuda.synchronize()
    return out_na.cpu().numpy(), out_ref.cpu().numpy()


def test_noalloc_canny_matches_the_ordinary_entry_point_twice(pf):
    import torch
    from libpillowfight_b200 import _capi
    if pf.device_count() <= 0:
        pytest.fail("no CUDA device: the gpu-marked tests must run on the B200 box")
    # first call: a large, busy page (many weak candidates land in the array)
    big = uniform_random_page(1500, 1100, seed=42)
    got, ref = _run_both(torch, _capi, big)
    assert np.array_equal(got, ref)
    assert (ref[..., 0] == 255).any()
    # second call, straight after: a smaller document page; the array still
    # holds the first call's list and links
    small = synthetic_page(620, 877, 7)
    got2, ref2 = _run_both(torch, _capi, small)
    assert np.array_equal(got2, ref2)
    assert (ref2[..., 0] == 255).any()
    # and it is the same answer the host-pointer entry point gives
    assert np.array_equal(got2, pf.canny(small))
    # ... and the CPU oracle, directly (not only through the other entry point)
    from oracle import pf_oracle
    assert np.array_equal(got, pf_oracle.canny(big))
    assert np.array_equal(got2, pf_oracle.canny(small))


def test_noalloc_canny_allocates_nothing(pf):
    """What the name promises: once the entry point has been called on the
    device (set-up), further calls leave the amount of free device memory where
    it was.  torch's caching allocator is kept out of the measured region (the
    buffers exist before it starts), so any change would be the library's."""
    import torch
    from libpillowfight_b200 import _capi
    if pf.device_count() <= 0:
        pytest.fail("no CUDA device: the gpu-marked tests must run on the B200 box")
    lib = _capi.lib
    dev = torch.device("cuda:0")
    page = synthetic_page(1024, 768, 3)
    h, w = page.shape[:2]
    tin = torch.from_numpy(page).to(dev)
    out = torch.zeros_like(tin)
    st = torch.cuda.current_stream().cuda_stream
    _capi.check(lib.pf_b200_dev_canny_noalloc(tin.data_ptr(), out.data_ptr(), w, h, st))   # set-up call
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info(dev)
    for _ in range(5):
        _capi.check(lib.pf_b200_dev_canny_noalloc(tin.data_ptr(), out.data_ptr(), w, h, st))
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info(dev)
    assert free_after == free_before
    from oracle import pf_oracle
    assert np.array_equal(out.cpu().numpy(), pf_oracle.canny(page))


def test_noalloc_canny_refuses_pages_beyond_its_limit(pf):
    import torch
    from libpillowfight_b200 import _capi
    if pf.device_count() <= 0:
        pytest.fail("no CUDA device: the gpu-marked tests must run on the B200 box")
    lib = _capi.lib
    dev = torch.device("cuda:0")
    w, h = 2480, 3508                       # A4 at 300 dpi: 8.7 Mpx > 4 Mpx
    assert w * h > MAX_PIXELS
    buf = torch.zeros((2, h, w, 4), dtype=torch.uint8, device=dev)
    rc = lib.pf_b200_dev_canny_noalloc(buf[0].data_ptr(), buf[1].data_ptr(), w, h, None)
    assert rc == _capi.PF_B200_ETOOBIG
    code, msg = _capi.last_error()
    assert code == _capi.PF_B200_ETOOBIG and "4194304" in msg
    # exactly at the limit is accepted
    w2, h2 = 2048, 2048
    assert w2 * h2 == MAX_PIXELS
    page = uniform_random_page(w2, h2, seed=1)
    got, ref = _run_both(torch, _capi, page)
    assert np.array_equal(got, ref)
