"""Can the kernels work on pinned host pages directly (no copies, PCIe traffic
issued by the SMs)?  Times the device entry points on host pointers."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from libpillowfight_b200 import _capi  # noqa: E402
from libpillowfight_b200.pages import A4_300DPI, synthetic_page  # noqa: E402

w, h = A4_300DPI
lib, chk = _capi.lib, _capi.check
st = torch.cuda.current_stream().cuda_stream
h_in = torch.from_numpy(synthetic_page(w, h, 0)).pin_memory()
h_out = torch.empty_like(h_in).pin_memory()
d_in = h_in.cuda()
d_out = torch.empty_like(d_in)
scratch = torch.empty(lib.pf_b200_canny_scratch_bytes(w, h), dtype=torch.uint8, device="cuda")


def timed(fn, n=10):
    for _ in range(2):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


for name, fn in (("sobel", lambda i, o: chk(lib.pf_b200_dev_sobel(i, o, w, h, st))),
                 ("gaussian", lambda i, o: chk(lib.pf_b200_dev_gaussian(i, o, w, h, 2.0, 5, st))),
                 ("canny", lambda i, o: chk(lib.pf_b200_dev_canny(i, o, w, h, 1, scratch.data_ptr(), scratch.numel(), st)))):
    t = [timed(lambda: fn(d_in.data_ptr(), d_out.data_ptr())),
         timed(lambda: fn(h_in.data_ptr(), d_out.data_ptr())),
         timed(lambda: fn(d_in.data_ptr(), h_out.data_ptr())),
         timed(lambda: fn(h_in.data_ptr(), h_out.data_ptr()))]
    ok = bool((h_out.cuda() == d_out).all())
    print("%-9s dev->dev %.3f  host->dev %.3f  dev->host %.3f  host->host %.3f ms   same result: %s" % ((name,) + tuple(t) + (ok,)))
