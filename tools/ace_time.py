"""Times the ACE accumulation kernel alone on the A4 page and reports the
mismatch rate against the oracle on a smaller page (0: the result is exact)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libpillowfight_b200 as pf  # noqa: E402
from libpillowfight_b200 import _capi  # noqa: E402
from libpillowfight_b200.distributed import xy_tensor  # noqa: E402
from libpillowfight_b200.pages import A4_300DPI, synthetic_page  # noqa: E402
from oracle import pf_oracle as o  # noqa: E402

w, h = A4_300DPI
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100
dev = torch.device("cuda:0")
lib, chk = _capi.lib, _capi.check
page = torch.from_numpy(synthetic_page(w, h, 0)).to(dev)
xy = xy_tensor(pf.ace_samples(n, w, h, 12345), dev)
scores = torch.empty(lib.pf_b200_ace_scores_bytes(w, h), dtype=torch.uint8, device=dev)
mm = torch.empty(lib.pf_b200_ace_minmax_bytes(w, h) // 8, dtype=torch.float64, device=dev)
st = torch.cuda.current_stream().cuda_stream


def run():
    chk(lib.pf_b200_dev_ace_scores(page.data_ptr(), w, h, 0, h, xy.data_ptr(), n, 10.0, 1000.0,
                                   scores.data_ptr(), mm.data_ptr(), st))


chk(lib.pf_b200_dev_ace_minmax_init(mm.data_ptr(), w, h, st))
for _ in range(3):
    run()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
a.record()
for _ in range(20):
    run()
b.record()
torch.cuda.synchronize()
ms = a.elapsed_time(b) / 20

small = synthetic_page(1240, 1754, 1)
ref = o.ace_with_samples(small, o.ace_draw_samples(n, 1240, 1754, 7), 10.0, 1000.0,
                         nb_threads=os.cpu_count())
got = pf.ace(small, 10, 1000, n, seed=7)
d = np.abs(got.astype(np.int16) - ref.astype(np.int16))[..., :3]
print("samples=%d: %.4f ms  (%.0f Mpx/s)  mismatch %.6f %% max %d" % (
    n, ms, w * h / ms / 1e3, 100.0 * (d != 0).mean(), d.max()))
