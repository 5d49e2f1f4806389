"""Canny on cold A4 pages, per call; also on a page of noise (long work list).

    python tools/canny_time.py
(a build with -DPF_CANNY_CLOCKS and PF_CANNY_CLOCKS=1 also prints the cycle
stamps of the short-list resolve)
"""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import libpillowfight_b200 as pf
from libpillowfight_b200 import _capi
from libpillowfight_b200.pages import A4_300DPI, synthetic_page
lib, chk = _capi.lib, _capi.check
w, h = A4_300DPI
dev = torch.device("cuda:0")
N = 12
st = torch.cuda.current_stream().cuda_stream
scratch = torch.empty(lib.pf_b200_canny_scratch_bytes(w, h), dtype=torch.uint8, device=dev)


def time_pages(pages, label):
    outs = torch.empty_like(pages)

    def run(i):
        chk(lib.pf_b200_dev_canny(pages[i % N].data_ptr(), outs[i % N].data_ptr(), w, h, 1,
                                  scratch.data_ptr(), scratch.numel(), st))
    for i in range(5):
        run(i)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for rep in range(3):
        torch.cuda.synchronize()
        a.record()
        for i in range(96):
            run(i)
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b) / 96 * 1000)
    weak = int(scratch[:16].view(torch.int32)[3].item())
    print("canny %-8s: %.2f us per call, %d weak pixels on the last page" % (label, best, weak), flush=True)


base = [synthetic_page(w, h, i) for i in range(2)]
time_pages(torch.stack([torch.from_numpy(np.roll(base[i % 2], 37 * i, axis=0)) for i in range(N)]).to(dev),
           "document")
rng = np.random.default_rng(5)
noise = rng.integers(0, 256, size=(2, h, w, 4), dtype=np.uint8)
time_pages(torch.stack([torch.from_numpy(np.roll(noise[i % 2], 37 * i, axis=0)) for i in range(N)]).to(dev),
           "noise")
if os.environ.get("PF_CANNY_CLOCKS"):
    pages = torch.stack([torch.from_numpy(base[0])]).to(dev)
    out = torch.empty_like(pages)
    for rep in range(3):
        chk(lib.pf_b200_dev_canny(pages[0].data_ptr(), out[0].data_ptr(), w, h, 1,
                                  scratch.data_ptr(), scratch.numel(), st))
        torch.cuda.synchronize()
        s = scratch[128:128 + 64].view(torch.int32).tolist()
        print("stamps (cycles): load %d, neighbours %d, rounds %d, write %d" % (
            s[1] - s[0], s[2] - s[1], s[3] - s[2], s[4] - s[3]))
