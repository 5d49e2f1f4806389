"""Runs each stage of the hot path a few times on one A4 page -- the short
program the ncu captures in profiles/ are taken from (see profiles/README.md)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libpillowfight_b200 as pf  # noqa: E402
from libpillowfight_b200 import _capi  # noqa: E402
from libpillowfight_b200.distributed import xy_tensor  # noqa: E402
from libpillowfight_b200.pages import A4_300DPI, synthetic_page  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
w, h = A4_300DPI
dev = torch.device("cuda:0")
lib, chk = _capi.lib, _capi.check
page = torch.from_numpy(synthetic_page(w, h, 0)).to(dev)
out = torch.empty_like(page)
out2 = torch.empty_like(page)
xy = xy_tensor(pf.ace_samples(100, w, h, 12345), dev)
scores = torch.empty(lib.pf_b200_ace_scores_bytes(w, h), dtype=torch.uint8, device=dev)
mm = torch.empty(lib.pf_b200_ace_minmax_bytes(w, h) // 8, dtype=torch.float64, device=dev)
scratch = torch.empty(lib.pf_b200_canny_scratch_bytes(w, h), dtype=torch.uint8, device=dev)
st = torch.cuda.current_stream().cuda_stream
for _ in range(reps):
    chk(lib.pf_b200_dev_ace_minmax_init(mm.data_ptr(), w, h, st))
    chk(lib.pf_b200_dev_ace_scores(page.data_ptr(), w, h, 0, h, xy.data_ptr(), 100, 10.0, 1000.0,
                                   scores.data_ptr(), mm.data_ptr(), st))
    chk(lib.pf_b200_dev_ace_refine(page.data_ptr(), w, h, 0, h, xy.data_ptr(), 100, 10.0, 1000.0,
                                   scores.data_ptr(), mm.data_ptr(), st))
    chk(lib.pf_b200_dev_ace_scale(page.data_ptr(), w, h, 0, h, xy.data_ptr(), 100, 10.0, 1000.0,
                                  scores.data_ptr(), mm.data_ptr(), out.data_ptr(), st))
    chk(lib.pf_b200_dev_gaussian(page.data_ptr(), out.data_ptr(), w, h, 2.0, 5, st))
    chk(lib.pf_b200_dev_sobel(out.data_ptr(), out2.data_ptr(), w, h, st))
    chk(lib.pf_b200_dev_canny(page.data_ptr(), out.data_ptr(), w, h, 1, scratch.data_ptr(),
                              scratch.numel(), st))
torch.cuda.synchronize()
print("ok")
