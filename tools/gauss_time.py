"""Times the Gaussian device entry point on three kinds of page (one library per process)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from libpillowfight_b200 import _capi  # noqa: E402
from libpillowfight_b200.pages import A4_300DPI, synthetic_page  # noqa: E402

w, h = A4_300DPI
dev = torch.device("cuda:0")
lib, chk = _capi.lib, _capi.check
st = torch.cuda.current_stream().cuda_stream
n = int(sys.argv[1]) if len(sys.argv) > 1 else 5
sigma = float(sys.argv[2]) if len(sys.argv) > 2 else 2.0
pages = {"synthetic": synthetic_page(w, h, 0), "white": np.full((h, w, 4), 255, dtype=np.uint8)}
out = torch.empty((h, w, 4), dtype=torch.uint8, device=dev)
res = []
for name, arr in pages.items():
    page = torch.from_numpy(arr).to(dev)
    for _ in range(3):
        chk(lib.pf_b200_dev_gaussian(page.data_ptr(), out.data_ptr(), w, h, sigma, n, st))
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(50):
        chk(lib.pf_b200_dev_gaussian(page.data_ptr(), out.data_ptr(), w, h, sigma, n, st))
    b.record()
    torch.cuda.synchronize()
    res.append("%s %.1f us" % (name, a.elapsed_time(b) / 50 * 1e3))
print("gaussian(%g, %d): " % (sigma, n) + ", ".join(res))
