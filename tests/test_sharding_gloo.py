"""World-size-2 gloo tests (CPU) of the multi-GPU host logic in
libpillowfight_b200/distributed.py.  The arithmetic is supplied by the oracle
here; on GPUs the same sharding code drives the C-ABI device entry points."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from libpillowfight_b200.distributed import ace_sharded, page_shard, row_bands
from libpillowfight_b200.pages import synthetic_page


class OracleAceCompute:
    """Same interface as CudaAceCompute, CPU tensors, oracle arithmetic."""

    def __init__(self):
        from oracle import pf_oracle
        self.o = pf_oracle

    def new_minmax(self, device, width, height):
        big = np.finfo(np.float64).max
        return torch.tensor(([big] * 3 + [-big] * 3) * 2, dtype=torch.float64)

    def refine(self, page, xy, slope, limit, band, scores, minmax):
        return False        # the oracle's scores are the reference's already

    def sample_pixels(self, page, xy):
        p, c = page.numpy(), xy.numpy()
        return torch.from_numpy(np.ascontiguousarray(p[c[1::2], c[0::2]]))

    def scores_band(self, rows, height, band, xy, spx, slope, limit, minmax):
        # the oracle wants a page: one that is right in the band's rows and at
        # the samples' coordinates, and wrong (0x77) everywhere else
        b0, b1 = band
        c = xy.numpy()
        fake = np.full((height, rows.shape[1], 4), 0x77, dtype=np.uint8)
        fake[c[1::2], c[0::2]] = spx.numpy()
        fake[b0:b1] = rows.numpy()
        return self.scores(torch.from_numpy(fake), xy, slope, limit, band, minmax)

    def refine_band(self, rows, height, band, xy, spx, slope, limit, scores, minmax):
        return False

    def scores(self, page, xy, slope, limit, band, minmax):
        b0, b1 = band
        s = self.o.ace_scores(page.numpy(), xy.numpy(), slope, limit, b0, b1)[:, b0:b1]
        for c in range(3):
            minmax[c] = min(float(minmax[c]), float(np.nanmin(s[c])))
            minmax[3 + c] = max(float(minmax[3 + c]), float(np.nanmax(s[c])))
        return torch.from_numpy(np.ascontiguousarray(s))

    def scale(self, page, xy, slope, limit, band, scores, minmax):
        return self._scale(scores, minmax)

    def scale_band(self, rows, height, band, xy, spx, slope, limit, scores, minmax):
        return self._scale(scores, minmax)

    def _scale(self, scores, minmax):
        s = scores.numpy()
        mn = minmax[:3].numpy()[:, None, None]
        mx = minmax[3:6].numpy()[:, None, None]
        with np.errstate(invalid="ignore"):
            v = 255.0 * (s - mn) / (mx - mn)
        v = np.where(v > 0.0, np.minimum(v, 255.0), 0.0)
        out = np.empty(s.shape[1:] + (4,), dtype=np.uint8)
        out[..., :3] = np.moveaxis(v.astype(np.int32).astype(np.uint8), 0, -1)
        out[..., 3] = 255
        return torch.from_numpy(out)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, w, h, n_samples, result_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import pf_oracle as o
        if rank == 0:
            page = torch.from_numpy(synthetic_page(w, h, 5))
            xy = torch.from_numpy(o.ace_draw_samples(n_samples, w, h, 99))
        else:
            page = torch.zeros((h, w, 4), dtype=torch.uint8)
            xy = torch.zeros(2 * n_samples, dtype=torch.int32)
        out = ace_sharded(page, xy, 10.0, 1000.0, OracleAceCompute())
        # (only rank 0 ever held the page: the others were sent their band and
        # the samples' pixel values, nothing else)
        whole = o.ace_with_samples(synthetic_page(w, h, 5), xy.numpy(), 10.0, 1000.0)
        assert np.array_equal(out.numpy(), whole), "rank %d: sharded != whole" % rank
        if rank == 0:
            np.save(result_path, out.numpy())
    finally:
        dist.destroy_process_group()


def test_row_bands_cover_the_page():
    for h in (1, 7, 3508, 7016):
        for world in (1, 2, 3, 4, 8):
            if world > h:
                continue
            bands = row_bands(h, world)
            assert bands[0][0] == 0 and bands[-1][1] == h
            for (a0, a1), (b0, b1) in zip(bands, bands[1:]):
                assert a1 == b0
            sizes = [b - a for a, b in bands]
            assert max(sizes) - min(sizes) <= 1


def test_page_shards_partition_the_batch():
    for n in (1, 5, 512):
        for world in (1, 2, 4, 8):
            seen = sorted(i for r in range(world) for i in page_shard(n, world, r))
            assert seen == list(range(n))


import pytest


@pytest.mark.parametrize("h", [75, 76])     # bands of 37 + 38 rows (padded gather), and 38 + 38
def test_ace_row_band_sharding_world2(tmp_path, oracle, h):
    w, n = 64, 12
    port = _free_port()
    res = str(tmp_path / "out.npy")
    mp.spawn(_worker, args=(2, port, w, h, n, res), nprocs=2, join=True)
    out = np.load(res)
    assert out.shape == (h, w, 4)
