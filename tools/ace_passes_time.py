"""Times the ACE passes one by one in the order a page goes through them."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libpillowfight_b200 as pf  # noqa: E402
from libpillowfight_b200 import _capi  # noqa: E402
from libpillowfight_b200.distributed import xy_tensor  # noqa: E402
from libpillowfight_b200.pages import A4_300DPI, synthetic_page  # noqa: E402

w, h = A4_300DPI
dev = torch.device("cuda:0")
lib, chk = _capi.lib, _capi.check
pages = [torch.from_numpy(synthetic_page(w, h, i)).to(dev) for i in range(4)]
out = torch.empty_like(pages[0])
xy = xy_tensor(pf.ace_samples(100, w, h, 12345), dev)
scores = torch.empty(lib.pf_b200_ace_scores_bytes(w, h), dtype=torch.uint8, device=dev)
mm = torch.empty(lib.pf_b200_ace_minmax_bytes(w, h) // 8, dtype=torch.float64, device=dev)
st = torch.cuda.current_stream().cuda_stream
args = lambda p: (p.data_ptr(), w, h, 0, h, xy.data_ptr(), 100, 10.0, 1000.0, scores.data_ptr(), mm.data_ptr(), st)
names = ["init", "scores", "refine", "scale+fixup"]
tot = np.zeros(4)
reps = 12
for it in range(reps + 3):
    p = pages[it % 4]
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
    ev[0].record()
    chk(lib.pf_b200_dev_ace_minmax_init(mm.data_ptr(), w, h, st)); ev[1].record()
    chk(lib.pf_b200_dev_ace_scores(*args(p))); ev[2].record()
    chk(lib.pf_b200_dev_ace_refine(*args(p))); ev[3].record()
    chk(lib.pf_b200_dev_ace_scale(*args(p)[:-1], out.data_ptr(), st)); ev[4].record()
    torch.cuda.synchronize()
    if it >= 3:
        tot += [ev[i].elapsed_time(ev[i + 1]) for i in range(4)]
    if it == 3:
        m = mm.cpu().numpy()
        tiles = m[12:].reshape(-1, 6)
        thr = 2 * 8e-6
        wanted = ((tiles[:, :3] <= m[0:3] + thr) | (tiles[:, 3:] >= m[3:6] - thr)).any(1)
        print("fast min/max", m[:6], "exact", m[6:12], "tiles wanted", int(wanted.sum()), "of", len(tiles))
print("pixels redone by the fix-up pass: %d of %d" % (int(scores[:256].view(torch.int32)[32]), w * h))
for n, t in zip(names, tot / reps):
    print("%-12s %.4f ms" % (n, t))
