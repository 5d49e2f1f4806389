import os, sys, torch
sys.path.insert(0, '/root/repo')
import libpillowfight_b200 as pf
from libpillowfight_b200 import _capi
from libpillowfight_b200.pages import A4_300DPI, synthetic_page
lib, chk = _capi.lib, _capi.check
w, h = A4_300DPI
dev = torch.device("cuda:0")
pages = torch.stack([torch.from_numpy(synthetic_page(w, h, i)) for i in range(4)]).to(dev)
outs = torch.empty_like(pages)
scratch = torch.empty(lib.pf_b200_canny_scratch_bytes(w, h), dtype=torch.uint8, device=dev)
st = torch.cuda.current_stream().cuda_stream
def run(i):
    chk(lib.pf_b200_dev_canny(pages[i % 4].data_ptr(), outs[i % 4].data_ptr(), w, h, 1, scratch.data_ptr(), scratch.numel(), st))
for i in range(5): run(i)
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); a.record()
for i in range(100): run(i)
b.record(); torch.cuda.synchronize()
print("canny debug=%s: %.2f us per call" % (os.environ.get("PF_B200_CANNY_DEBUG", "0"), a.elapsed_time(b) * 10))
