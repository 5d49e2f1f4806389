"""Per-phase timeline of the row-band ACE over NVLink peer memory
(BASELINE.json config 4: 4960 x 7016, 400 samples).  Run under torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \\
        tools/ace_bands_timeline.py

Prints, per phase, the time between the CUDA events that bracket it: the
maximum and the minimum over the ranks, averaged over the repetitions."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libpillowfight_b200 as pf  # noqa: E402
from libpillowfight_b200.distributed import AcePeerBands, xy_tensor  # noqa: E402
from libpillowfight_b200.pages import A4_300DPI, A4_600DPI, synthetic_page  # noqa: E402

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
rank, world = dist.get_rank(), dist.get_world_size()
w, h = A4_600DPI
peer = AcePeerBands(h, w, dev)
if rank == 0:
    small = torch.from_numpy(synthetic_page(*A4_300DPI, page_index=0)).to(dev)
    peer.page.copy_(small.repeat_interleave(2, dim=0).repeat_interleave(2, dim=1)[:h, :w])
xy = xy_tensor(pf.ace_samples(400, w, h, 12345), dev)
for _ in range(3):
    peer.run(xy, 10.0, 1000.0)
reps = 10
acc = None
for _ in range(reps):
    dist.barrier()
    torch.cuda.synchronize()
    marks = []
    peer.run(xy, 10.0, 1000.0, marks=marks)
    torch.cuda.synchronize()
    ms = np.array([marks[i][1].elapsed_time(marks[i + 1][1]) for i in range(len(marks) - 1)])
    acc = ms if acc is None else acc + ms
names = [m[0] for m in marks[1:]]
t = torch.tensor(acc / reps, dtype=torch.float64, device=dev)
tmax, tmin = t.clone(), t.clone()
dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
if rank == 0:
    print("ACE 4960x7016, 400 samples, row bands over %d GPUs (peer memory, %s): ms per phase" % (
        world, "multimem.st" if peer.use_multicast else "per-peer stores"))
    for n, a, b in zip(names, tmax.tolist(), tmin.tolist()):
        print("  %-44s max %.3f  min %.3f" % (n, a, b))
    print("  %-44s %.3f (sum of the maxima)" % ("total", float(tmax.sum())))
dist.destroy_process_group()
