"""Times the device entry points on pages of different content (the Gaussian's
exact-order fallback and Canny's hysteresis depend on it)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libpillowfight_b200 as pf  # noqa: E402
from libpillowfight_b200 import _capi  # noqa: E402
from libpillowfight_b200.distributed import xy_tensor  # noqa: E402
from libpillowfight_b200.pages import A4_300DPI, synthetic_page, uniform_random_page  # noqa: E402

w, h = A4_300DPI
dev = torch.device("cuda:0")
lib, chk = _capi.lib, _capi.check
st = torch.cuda.current_stream().cuda_stream
rng = np.random.default_rng(0)
white = np.full((h, w, 4), 255, dtype=np.uint8)
text = white.copy()                       # black 3x12 strokes on white, ~4 % coverage
for _ in range(28000):
    y, x = rng.integers(0, h - 12), rng.integers(0, w - 3)
    text[y:y + 12, x:x + 3, :3] = 0
gray = np.full((h, w, 4), 255, dtype=np.uint8)
gray[..., :3] = 137
# a smooth texture (blurred noise, stretched): many weak edges, like a photograph
tex = rng.normal(size=(h // 4 + 2, w // 4 + 2))
tex = np.kron(tex, np.ones((4, 4)))[:h, :w]
for _ in range(3):
    tex = (tex + np.roll(tex, 1, 0) + np.roll(tex, -1, 0) + np.roll(tex, 1, 1) + np.roll(tex, -1, 1)) / 5
tex = np.clip(128 + tex * 160, 0, 255).astype(np.uint8)
texture = np.stack([tex, tex, tex, np.full_like(tex, 255)], -1)
blank = np.full((h, w, 4), 255, dtype=np.uint8)
blank[..., :3] = np.clip(232 + rng.integers(-2, 3, size=(h, w, 3)), 0, 255).astype(np.uint8)
contents = {"blank + noise 2": blank, "texture": np.ascontiguousarray(texture), "synthetic": synthetic_page(w, h, 0), "noise": uniform_random_page(w, h, seed=1),
            "white": white, "flat gray": gray, "text on white": text}
out = torch.empty((h, w, 4), dtype=torch.uint8, device=dev)
scratch = torch.empty(lib.pf_b200_canny_scratch_bytes(w, h), dtype=torch.uint8, device=dev)
scores = torch.empty(lib.pf_b200_ace_scores_bytes(w, h), dtype=torch.uint8, device=dev)
mm = torch.empty(lib.pf_b200_ace_minmax_bytes(w, h) // 8, dtype=torch.float64, device=dev)
xy = xy_tensor(pf.ace_samples(100, w, h, 12345), dev)


def timed(fn, n=20):
    for _ in range(3):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3


print("%-14s %9s %9s %9s %9s %9s %9s  (us)" % ("content", "gaussian", "sobel", "canny", "ace", "refine", "scale+fix"))
for name, arr in contents.items():
    page = torch.from_numpy(arr).to(dev)
    ace_args = (page.data_ptr(), w, h, 0, h, xy.data_ptr(), 100, 10.0, 1000.0, scores.data_ptr(), mm.data_ptr(), st)

    def ace_all():
        chk(lib.pf_b200_dev_ace_minmax_init(mm.data_ptr(), w, h, st))
        chk(lib.pf_b200_dev_ace_scores(*ace_args))

    ace_all()
    t = [timed(lambda: chk(lib.pf_b200_dev_gaussian(page.data_ptr(), out.data_ptr(), w, h, 2.0, 5, st))),
         timed(lambda: chk(lib.pf_b200_dev_sobel(page.data_ptr(), out.data_ptr(), w, h, st))),
         timed(lambda: chk(lib.pf_b200_dev_canny(page.data_ptr(), out.data_ptr(), w, h, 1, scratch.data_ptr(),
                                                 scratch.numel(), st))),
         timed(ace_all),
         timed(lambda: chk(lib.pf_b200_dev_ace_refine(*ace_args))),
         timed(lambda: chk(lib.pf_b200_dev_ace_scale(*ace_args[:-1], out.data_ptr(), st)))]
    import ctypes
    hdr = torch.zeros(4, dtype=torch.int32)
    hdr.copy_(scratch[:16].view(torch.int32))
    print("%-14s %9.1f %9.1f %9.1f %9.1f %9.1f %9.1f   weak pixels %d  ace pixels redone %d" % (
        (name,) + tuple(t) + (int(hdr[3]), int(scores[:256].view(torch.int32)[32]))))
