"""Multi-GPU sharding of the hot path: one process per GPU, torch.distributed
for the plumbing (NCCL on GPUs, gloo in the CPU tests).

Only the two splits BASELINE.json names:
  * ACE: every rank gets its band of rows (scatter) and the N sample pixels'
    values (a few hundred bytes, broadcast) -- not the page: a band's scores
    need nothing else.  Each rank accumulates the scores of its band, the
    per-channel min/max are all-reduced (6 doubles -- the one real exchange
    the algorithm has; a second time for the scores at the extremes once
    they have been recomputed exactly), each rank rescales its band, and the
    bands are all-gathered.
  * Gaussian/Sobel/Canny: pages are independent; page i goes to rank
    i mod world and nothing is exchanged.

The arithmetic is behind a small `compute` object so that the same sharding
code runs in the CPU gloo tests (tests/ plug the oracle in there); the product
object, CudaAceCompute, calls the C-ABI device entry points on torch CUDA
tensors' data pointers.
"""
import numpy as np
import torch
import torch.distributed as dist


def row_bands(height, world):
    """[begin, end) rows of each rank; sizes differ by at most one row."""
    return [(height * r // world, height * (r + 1) // world) for r in range(world)]


def page_shard(n_pages, world, rank):
    """Indices of the pages rank `rank` processes (round robin)."""
    return list(range(rank, n_pages, world))


class CudaAceCompute:
    """ACE band arithmetic on the current CUDA device through the C-ABI.

    A band's `scores` is its score block (include/pillowfight_b200.h): an
    opaque uint8 CUDA tensor of pf_b200_ace_scores_bytes(width, rows) bytes."""

    def __init__(self):
        from . import _capi
        self._capi = _capi
        self._lib = _capi.lib

    def new_minmax(self, device, width, height):
        """The min/max block of one page (include/pillowfight_b200.h): [0:3]
        min, [3:6] max of the approximate scores, [6:9], [9:12] the exact ones."""
        n = self._lib.pf_b200_ace_minmax_bytes(width, height) // 8
        mm = torch.empty(n, dtype=torch.float64, device=device)
        st = torch.cuda.current_stream().cuda_stream
        self._capi.check(self._lib.pf_b200_dev_ace_minmax_init(mm.data_ptr(), width, height, st))
        return mm

    def new_scores(self, device, width, rows):
        return torch.empty(self._lib.pf_b200_ace_scores_bytes(width, rows), dtype=torch.uint8,
                           device=device)

    @staticmethod
    def score_planes(scores, width, rows):
        """The three FP32 score planes inside a score block, (3, rows, W)."""
        return scores[256:256 + 12 * width * rows].view(torch.float32).view(3, rows, width)

    def scores(self, page, xy, slope, limit, band, minmax, out=None):
        """page: (H, W, 4) uint8 CUDA tensor; xy: int32 CUDA tensor (2N);
        returns the band's score block, folds min/max into `minmax`."""
        h, w = page.shape[0], page.shape[1]
        b0, b1 = band
        out = out if out is not None else self.new_scores(page.device, w, b1 - b0)
        st = torch.cuda.current_stream().cuda_stream
        self._capi.check(self._lib.pf_b200_dev_ace_scores(
            page.data_ptr(), w, h, b0, b1, xy.data_ptr(), xy.numel() // 2,
            float(slope), float(limit), out.data_ptr(), minmax.data_ptr(), st))
        return out

    def refine(self, page, xy, slope, limit, band, scores, minmax):
        """The band's part of the exact extremes, into minmax[6:12].  True:
        those need the exchange too."""
        h, w = page.shape[0], page.shape[1]
        st = torch.cuda.current_stream().cuda_stream
        self._capi.check(self._lib.pf_b200_dev_ace_refine(
            page.data_ptr(), w, h, band[0], band[1], xy.data_ptr(), xy.numel() // 2,
            float(slope), float(limit), scores.data_ptr(), minmax.data_ptr(), st))
        return True

    def sample_pixels(self, page, xy):
        """The samples' pixel values, (N, 4) uint8, where the whole page is."""
        px = page.view(-1, 4)
        idx = xy[1::2].long() * page.shape[1] + xy[0::2].long()
        return px.index_select(0, idx).contiguous()

    def scores_band(self, rows, height, band, xy, spx, slope, limit, minmax, out=None):
        """rows: (b1-b0, W, 4) uint8 CUDA tensor holding only the band."""
        b0, b1 = band
        w = rows.shape[1]
        out = out if out is not None else self.new_scores(rows.device, w, b1 - b0)
        st = torch.cuda.current_stream().cuda_stream
        self._capi.check(self._lib.pf_b200_dev_ace_scores_band(
            rows.data_ptr(), w, height, b0, b1, xy.data_ptr(), spx.data_ptr(), xy.numel() // 2,
            float(slope), float(limit), out.data_ptr(), minmax.data_ptr(), st))
        return out

    def refine_band(self, rows, height, band, xy, spx, slope, limit, scores, minmax):
        w = rows.shape[1]
        st = torch.cuda.current_stream().cuda_stream
        self._capi.check(self._lib.pf_b200_dev_ace_refine_band(
            rows.data_ptr(), w, height, band[0], band[1], xy.data_ptr(), spx.data_ptr(),
            xy.numel() // 2, float(slope), float(limit), scores.data_ptr(), minmax.data_ptr(), st))
        return True

    def scale(self, page, xy, slope, limit, band, scores, minmax):
        """The band's rows of the result, (rows, W, 4) uint8."""
        h, w = page.shape[0], page.shape[1]
        out = torch.empty((band[1] - band[0], w, 4), dtype=torch.uint8, device=page.device)
        st = torch.cuda.current_stream().cuda_stream
        self._capi.check(self._lib.pf_b200_dev_ace_scale(
            page.data_ptr(), w, h, band[0], band[1], xy.data_ptr(), xy.numel() // 2,
            float(slope), float(limit), scores.data_ptr(), minmax.data_ptr(), out.data_ptr(), st))
        return out

    def scale_band(self, rows, height, band, xy, spx, slope, limit, scores, minmax):
        w = rows.shape[1]
        out = torch.empty((band[1] - band[0], w, 4), dtype=torch.uint8, device=rows.device)
        st = torch.cuda.current_stream().cuda_stream
        self._capi.check(self._lib.pf_b200_dev_ace_scale_band(
            rows.data_ptr(), w, height, band[0], band[1], xy.data_ptr(), spx.data_ptr(),
            xy.numel() // 2, float(slope), float(limit), scores.data_ptr(), minmax.data_ptr(),
            out.data_ptr(), st))
        return out


def ace_sharded(page, xy, slope, limit, compute, group=None, src=0):
    """Row-band sharded ACE.  `page` must hold the input on rank `src`; on the
    other ranks only its shape and device are used.  `xy` is the sample table
    on rank `src` and a buffer of the same size elsewhere.  Every rank returns
    the whole result."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    h, w = page.shape[0], page.shape[1]
    minmax = compute.new_minmax(page.device, w, h)
    if world == 1:
        scores = compute.scores(page, xy, slope, limit, (0, h), minmax)
        compute.refine(page, xy, slope, limit, (0, h), scores, minmax)
        return compute.scale(page, xy, slope, limit, (0, h), scores, minmax)

    bands = row_bands(h, world)
    max_rows = max(b1 - b0 for b0, b1 in bands)
    b0, b1 = bands[rank]
    # the sample table and the samples' pixel values go to everyone; the page
    # goes out in bands (all of one size for the scatter: short ones padded)
    n = xy.numel() // 2
    spx = compute.sample_pixels(page, xy) if rank == src else \
        torch.empty((n, 4), dtype=torch.uint8, device=page.device)
    dist.broadcast(xy, src=src, group=group)
    dist.broadcast(spx, src=src, group=group)
    mine = torch.empty((max_rows, w, 4), dtype=torch.uint8, device=page.device)
    chunks = None
    if rank == src:
        chunks = []
        for r0, r1 in bands:
            if r1 - r0 == max_rows:
                chunks.append(page[r0:r1])
            else:
                pad = torch.zeros((max_rows, w, 4), dtype=torch.uint8, device=page.device)
                pad[: r1 - r0] = page[r0:r1]
                chunks.append(pad)
    dist.scatter(mine, scatter_list=chunks, src=src, group=group)
    rows = mine[: b1 - b0]

    scores = compute.scores_band(rows, h, (b0, b1), xy, spx, slope, limit, minmax)
    # The one exchange the algorithm has: the page's extremes.  A band finds the
    # exact extremes of its own rows from its own approximate ones (refine), so
    # only the exact ones travel.  (A compute object without a refinement pass
    # -- the CPU tests' oracle -- has exact scores already: its first six
    # doubles are exchanged instead.)
    if not compute.refine_band(rows, h, (b0, b1), xy, spx, slope, limit, scores, minmax):
        dist.all_reduce(minmax[0:3], op=dist.ReduceOp.MIN, group=group)
        dist.all_reduce(minmax[3:6], op=dist.ReduceOp.MAX, group=group)
    else:
        dist.all_reduce(minmax[6:9], op=dist.ReduceOp.MIN, group=group)
        dist.all_reduce(minmax[9:12], op=dist.ReduceOp.MAX, group=group)
    band = compute.scale_band(rows, h, (b0, b1), xy, spx, slope, limit, scores, minmax)
    out = torch.empty((h, w, 4), dtype=torch.uint8, device=page.device)
    if h % world == 0:
        # equal bands: gather straight into the result
        dist.all_gather_into_tensor(out.view(-1), band.contiguous().view(-1), group=group)
        return out
    # all-gather needs equal shapes: pad every band to the tallest
    padded = torch.zeros((max_rows, w, 4), dtype=torch.uint8, device=page.device)
    padded[: band.shape[0]] = band
    gathered = torch.empty((world, max_rows, w, 4), dtype=torch.uint8, device=page.device)
    dist.all_gather_into_tensor(gathered.view(-1), padded.view(-1), group=group)
    for r, (r0, r1) in enumerate(bands):
        out[r0:r1] = gathered[r, : r1 - r0]
    return out


class AcePeerBands:
    """Row-band ACE with the page and the result in NVLink peer memory (torch
    symmetric memory): no scatter and no all-gather.  Every GPU reads its band
    of rows straight out of the page on the GPU that holds it (the
    accumulation kernel loads each pixel once), and the rescale kernel stores
    its band into every GPU's result (pf_b200_dev_ace_scale_to).  What is left
    for NCCL is the sample table, the samples' pixel values and the two
    6-double min/max exchanges.

    One object per page size; `page` (on rank `src`) is where the caller puts
    the input, `out` holds the whole result on every rank after run()."""

    def __init__(self, height, width, device, group=None, use_multicast=True):
        import torch.distributed._symmetric_memory as symm
        from . import _capi
        self._capi, self._lib = _capi, _capi.lib
        self.h, self.w = height, width
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        self.page = symm.empty((height, width, 4), dtype=torch.uint8, device=device)
        self.out = symm.empty((height, width, 4), dtype=torch.uint8, device=device)
        self._hp = symm.rendezvous(self.page, self.group)
        self._ho = symm.rendezvous(self.out, self.group)
        self.compute = CudaAceCompute()
        # NVLS multicast for the result stores when the fabric offers it
        self.use_multicast = bool(use_multicast and getattr(self._ho, "multicast_ptr", 0))
        self._rows = self._dests = None
        self._tmp = torch.empty(6, dtype=torch.float64, device=device)
        self._msg = None
        # the sample table and the samples' pixel values travel through peer
        # memory too (up to this many samples; more go by NCCL broadcast)
        self.max_peer_samples = 4096
        self._msg_sym = symm.empty(3 * self.max_peer_samples, dtype=torch.int32, device=device)
        self._hm = symm.rendezvous(self._msg_sym, self.group)
        self._msg_src = {}
        self._minmax = torch.empty(self._lib.pf_b200_ace_minmax_bytes(width, height) // 8,
                                   dtype=torch.float64, device=device)
        b0, b1 = row_bands(height, self.world)[self.rank]
        self._scores = self.compute.new_scores(device, width, b1 - b0)

    def _exchange(self, minmax, lo):
        """min over minmax[lo:lo+3], max over minmax[lo+3:lo+6], in ONE
        all-reduce: -min is a max, and negation is exact."""
        t = self._tmp
        torch.neg(minmax[lo:lo + 3], out=t[0:3])
        t[3:6].copy_(minmax[lo + 3:lo + 6])
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        torch.neg(t[0:3], out=minmax[lo:lo + 3])
        minmax[lo + 3:lo + 6].copy_(t[3:6])

    def run(self, xy, slope, limit, src=0, marks=None):
        """marks: a list to which (phase name, CUDA event recorded at the END
        of that phase) pairs are appended (tools/ace_bands_timeline.py)."""
        def mark(name):
            if marks is not None:
                ev = torch.cuda.Event(enable_timing=True)
                ev.record()
                marks.append((name, ev))

        mark("start")
        h, w, comp = self.h, self.w, self.compute
        b0, b1 = row_bands(h, self.world)[self.rank]
        n = xy.numel() // 2
        dev = self.page.device
        # sample table and sample values in one message; it also orders every
        # rank's reads of the page after the copy that put it there (same
        # stream on `src`)
        if self._msg is None or self._msg.numel() != 3 * n:
            self._msg = torch.empty(3 * n, dtype=torch.int32, device=dev)
        msg = self._msg
        by_peer = n <= self.max_peer_samples
        if self.rank == src:
            if by_peer:
                msg = self._msg_sym[: 3 * n]
            self._capi.check(self._lib.pf_b200_dev_ace_gather_samples(
                self.page.data_ptr(), w, h, xy.data_ptr(), n, msg.data_ptr(),
                torch.cuda.current_stream().cuda_stream))
        if by_peer:
            # the holder's copy is in place: everyone else fetches it over NVLink
            # (a device-side barrier and one small copy instead of a broadcast)
            self._hm.barrier()
            if self.rank != src:
                if src not in self._msg_src:
                    self._msg_src[src] = self._hm.get_buffer(src, (3 * self.max_peer_samples,), torch.int32)
                msg.copy_(self._msg_src[src][: 3 * n])
        else:
            dist.broadcast(msg, src=src, group=self.group)
        mark("broadcast of the samples")
        xy, spx = msg[: 2 * n], msg[2 * n:]
        if self._rows is None:
            self._rows = self._hp.get_buffer(src, (h, w, 4), torch.uint8)[b0:b1]       # peer memory
        rows = self._rows
        st = torch.cuda.current_stream().cuda_stream
        minmax = self._minmax
        self._capi.check(self._lib.pf_b200_dev_ace_minmax_init(minmax.data_ptr(), w, h, st))
        scores = comp.scores_band(rows, h, (b0, b1), xy, spx, slope, limit, minmax, out=self._scores)
        mark("accumulation (band read over NVLink)")
        # each band finds the exact extremes of its own rows; those are the one
        # thing the bands have to tell each other
        comp.refine_band(rows, h, (b0, b1), xy, spx, slope, limit, scores, minmax)
        mark("exact extremes of the band")
        self._exchange(minmax, 6)
        mark("exchange of the extremes")
        mc = self._ho.multicast_ptr if self.use_multicast else 0
        args = (rows.data_ptr(), w, h, b0, b1, xy.data_ptr(), spx.data_ptr(), n, float(slope),
                float(limit), scores.data_ptr(), minmax.data_ptr())
        if mc:
            # one multimem.st per 16 bytes: the NVSwitch delivers the band to all GPUs
            self._capi.check(self._lib.pf_b200_dev_ace_scale_multicast(*args, mc + b0 * w * 4, st))
        else:
            import ctypes
            if self._dests is None:          # own image first, then the peers'
                order = [self.rank] + [r for r in range(self.world) if r != self.rank]
                self._dests = [self._ho.get_buffer(r, (h, w, 4), torch.uint8)[b0:b1] for r in order]
            ptrs = (ctypes.c_void_p * len(self._dests))(*[d.data_ptr() for d in self._dests])
            self._capi.check(self._lib.pf_b200_dev_ace_scale_to(*args, ptrs, len(self._dests), st))
        mark("rescale + fix-up, stores to every GPU")
        self._ho.barrier()          # every band has landed everywhere
        mark("barrier")
        return self.out


class CudaChain:
    """Gaussian -> Sobel -> Canny on batches of same-size pages resident in
    HBM, through the C-ABI device entry points."""

    def __init__(self):
        from . import _capi
        self._capi = _capi
        self._lib = _capi.lib
        self._scratch = None

    def _stream(self):
        return torch.cuda.current_stream().cuda_stream

    def gaussian(self, pages, out, sigma=2.0, nb_stddev=5):
        n, h, w = pages.shape[0], pages.shape[1], pages.shape[2]
        self._capi.check(self._lib.pf_b200_dev_gaussian_batch(
            pages.data_ptr(), out.data_ptr(), w, h, n, float(sigma), int(nb_stddev),
            self._stream()))
        return out

    def sobel(self, pages, out):
        n, h, w = pages.shape[0], pages.shape[1], pages.shape[2]
        self._capi.check(self._lib.pf_b200_dev_sobel_batch(
            pages.data_ptr(), out.data_ptr(), w, h, n, self._stream()))
        return out

    def canny(self, pages, out, pages_per_launch=1):
        n, h, w = pages.shape[0], pages.shape[1], pages.shape[2]
        need = 256 + 8 * w * h * max(1, min(n, pages_per_launch))
        if self._scratch is None or self._scratch.numel() < need \
                or self._scratch.device != pages.device:
            self._scratch = torch.empty(need, dtype=torch.uint8, device=pages.device)
        self._capi.check(self._lib.pf_b200_dev_canny(
            pages.data_ptr(), out.data_ptr(), w, h, n,
            self._scratch.data_ptr(), self._scratch.numel(), self._stream()))
        return out


def to_numpy_u8(t):
    return t.detach().cpu().numpy().astype(np.uint8, copy=False)


def xy_tensor(xy, device):
    return torch.from_numpy(np.ascontiguousarray(xy, dtype=np.int32)).to(device)
