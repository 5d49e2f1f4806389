#!/usr/bin/env python
"""Benchmark of the libpillowfight hot path on B200 (contract: task prompt, section 4).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One STEP = the whole hot path over one synthetic A4 300-dpi page per GPU:
    pf_apply_ace(slope 10, limit 1000, 100 samples, fixed seed)
    pf_apply_gaussian(sigma 2.0, nb_stddev 5) -> pf_apply_sobel on its output
    pf_apply_canny
metric = page pixels through that step per second (Mpx/s), whole job.
`value`: pages resident in HBM, C-ABI device entry points, CUDA-event timed.
`e2e`:   the host-pointer entry points (pf_b200_ace/gaussian/sobel/canny) on
         pinned host pages, copies inside the timed region.
At N > 1 every rank runs the step on its own pages (pages are independent:
no data-path collective, weak scaling); the row-band ACE split with its NCCL
broadcast / all-reduce / all-gather (BASELINE.json config 4) is measured as
`extras.ace_600dpi_row_bands`.
"""
import argparse
import ctypes
import json
import os
import socket
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

def load_pages_module():
    """libpillowfight_b200/pages.py by path: importing it through the package
    would load the product library, which the reference arm must not touch."""
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "pf_b200_pages", os.path.join(ROOT, "libpillowfight_b200", "pages.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


A4 = (2480, 3508)
A4_600 = (4960, 7016)
ACE_SEED = 12345
FLOPS_PAIR = 22.0      # reference flops per pixel-sample pair that is not skipped
FLOPS_SKIPPED = 5.0    # distance + test only
N_SETS = 8             # page sets rotated through, > L2 in total
METRIC = "ACE & Gauss+Sobel+Canny Mpx/s, A4@300dpi"
# the same for both arms (what the numbers are quoted on); per-arm details go elsewhere
CONFIG = {"workload": "A4@300dpi page: ace(slope 10, limit 1000, 100 samples, seed 12345)"
                      " + gaussian(2.0, 5)->sobel + canny",
          "width": A4[0], "height": A4[1], "samples": 100, "seed": ACE_SEED}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    return ap.parse_args()


# --------------------------------------------------------------------------
# helpers
# --------------------------------------------------------------------------

def count_pairs(w, h, xy):
    """(#pairs not skipped, #pairs skipped) for the dist < h/5 rule."""
    t2 = (h // 5) ** 2
    ys = np.arange(h, dtype=np.int64)
    skipped = 0
    for sx, sy in zip(xy[0::2], xy[1::2]):
        r2 = t2 - (ys - int(sy)) ** 2          # need dx^2 < r2
        ok = r2 > 0
        half = np.zeros(h, dtype=np.int64)
        half[ok] = np.ceil(np.sqrt(r2[ok].astype(np.float64))).astype(np.int64) - 1
        lo = np.maximum(int(sx) - half, 0)
        hi = np.minimum(int(sx) + half, w - 1)
        skipped += int(np.where(ok, np.maximum(hi - lo + 1, 0), 0).sum())
    total = w * h * (len(xy) // 2)
    return total - skipped, skipped


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the GPU is busy."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,utilization.gpu,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons, power = [], None, set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 8:
                continue
            try:
                clk, mx, util = float(r[0]), float(r[1]), float(r[3])
            except ValueError:
                continue
            smax = mx
            if util >= 10.0:
                sm.append(clk)
                try:
                    power.append(float(r[2]))
                except ValueError:
                    pass
                for n, v in zip(names, r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
        allclk = [float(r[0]) for r in self.rows if len(r) >= 8 and r[0].replace(".", "").isdigit()]
        return {"sm_mhz": float(np.median(sm)) if sm else (float(np.median(allclk)) if allclk else None),
                "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples_under_load": len(sm), "samples": len(self.rows),
                "power_w_max": max(power) if power else None}


def bind_to_gpu_cpus(index):
    """Run this rank on the CPUs next to its GPU (and so allocate its pinned
    pages on that NUMA node): matters for the end-to-end number when several
    ranks stream pages over PCIe at once."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * i + b for i, wd in enumerate(words) for b in range(64) if (wd >> b) & 1}
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        if cpus:
            os.sched_setaffinity(0, cpus)
        return sorted(cpus)
    except Exception:      # noqa: BLE001  (no NVML, containers without the call, ...)
        return None


def load_json(path):
    try:
        with open(path) as f:
            return json.load(f)
    except (OSError, ValueError):
        return None


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


# --------------------------------------------------------------------------
# CPU side: the oracle as the reported baseline / the reference arm
# --------------------------------------------------------------------------

def cpu_step(o, pages, xy, cores, rows=None):
    """The same step on host cores with the CPU oracle, on the top `rows` rows
    of each page (None = whole pages): ACE with `cores` pthreads page after
    page (the reference's own threading), then the stencil chain on up to
    `cores` threads, one page per thread (the reference's stencils are
    single-threaded).  Returns seconds."""
    h = pages[0].shape[0]
    whole = rows is None or rows >= h
    t0 = time.perf_counter()
    for p in pages:
        if whole:
            o.ace_with_samples(p, xy, 10.0, 1000.0, nb_threads=cores)
        else:       # the adjustment pass (all but ~1 % of ACE) on a band of the full page
            o.ace_scores(p, xy, 10.0, 1000.0, row_begin=0, row_end=rows, nb_threads=cores)

    def chain(p):
        q = p if whole else np.ascontiguousarray(p[:rows])
        g = o.gaussian(q, 2.0, 5)
        o.sobel(g)
        o.canny(q)

    threads = [threading.Thread(target=chain, args=(p,)) for p in pages]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    return time.perf_counter() - t0


def cpu_pages_per_step(cores):
    return max(1, min(cores, 4))


def cpu_sample_rows(o, pages, xy, cores, steps, budget_s):
    """Rows per page so that `steps` CPU steps fit in about budget_s seconds."""
    h = pages[0].shape[0]
    probe = 96
    t = cpu_step(o, pages, xy, cores, rows=probe)
    rows = int(budget_s / max(steps, 1) / (t / probe))
    return h if rows >= h else max(32, rows)


def run_reference(args):
    """--impl reference: the CPU implementation of the path on the host cores.
    The reference itself cannot be built (no sources in the mount), so this is
    the oracle port (oracle/pf_oracle.c), kind "port"."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import pf_oracle as o
    synthetic_page = load_pages_module().synthetic_page
    o.lib()
    cores = os.cpu_count() or 1
    npg = cpu_pages_per_step(cores)
    w, h = A4
    pages = [synthetic_page(w, h, i) for i in range(npg)]
    xy = o.ace_draw_samples(100, w, h, ACE_SEED)
    rows = cpu_sample_rows(o, pages, xy, cores, args.steps + args.warmup, 150.0)
    for _ in range(args.warmup):
        cpu_step(o, pages, xy, cores, rows)
    t = 0.0
    for _ in range(args.steps):
        t += cpu_step(o, pages, xy, cores, rows)
    per = t / max(args.steps, 1)
    val = npg * w * min(rows, h) / per / 1e6
    sample = ("%d A4 300-dpi page(s) per step, top %d of %d rows each: ACE with %d pthreads page "
              "after page, then Gaussian->Sobel and Canny one page per thread"
              % (npg, min(rows, h), h, cores))
    emit(json.dumps({
        "impl": "reference",
        "metric": METRIC, "value": val, "unit": "Mpx/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": per * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": dict(CONFIG),
        "details": {"pages_per_step": npg, "rows_per_page": min(rows, h)},
        "cpu_baseline": {"value": val, "unit": "Mpx/s", "cores": cores, "kind": "port",
                         "sample": sample},
        "e2e": {"value": val, "unit": "Mpx/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# --------------------------------------------------------------------------
# GPU side
# --------------------------------------------------------------------------

class Stage:
    ACE, GAUSS, SOBEL, CANNY = range(4)
    NAMES = ["ace", "gaussian", "sobel", "canny"]
    LAUNCHES = [6, 1, 1, 4]     # kernels per call (memsets not counted)


def run_b200(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun the way the driver does
        os.execvp(sys.executable, [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                                   "--nproc-per-node", str(args.gpus), "--master-addr", "127.0.0.1",
                                   "--master-port", str(free_port()), os.path.abspath(__file__)]
                  + sys.argv[1:])
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 path has no CPU fallback "
                         "(use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    bind_to_gpu_cpus(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    from __graft_entry__ import load_build_module
    if rank == 0:
        load_build_module().build()
    if world > 1:
        dist.barrier()
    import libpillowfight_b200 as pf
    from libpillowfight_b200 import _capi
    from libpillowfight_b200.distributed import CudaAceCompute, CudaChain, ace_sharded, xy_tensor
    from libpillowfight_b200.pages import synthetic_page
    lib = _capi.lib

    def barrier():
        if world > 1:
            dist.barrier()

    w, h = A4
    npix = w * h
    sampler = ClockSampler(local)

    # ---- inputs resident in HBM: N_SETS distinct pages per rank --------------
    host_pages = [synthetic_page(w, h, rank * 64 + i) for i in range(2)]
    pages = torch.empty((N_SETS, h, w, 4), dtype=torch.uint8, device=dev)
    for i in range(N_SETS):
        # distinct content per set: page i%2 rolled by i rows (cheap to make)
        pages[i] = torch.from_numpy(np.roll(host_pages[i % 2], 37 * i, axis=0)).to(dev)
    xy_host = pf.ace_samples(100, w, h, ACE_SEED)
    xy = xy_tensor(xy_host, dev)
    out_ace = torch.empty_like(pages)
    out_g = torch.empty_like(pages)
    out_s = torch.empty_like(pages)
    out_c = torch.empty_like(pages)
    st = torch.cuda.current_stream().cuda_stream
    chk = _capi.check

    class Slot:
        """Work areas of one page in flight, and the two streams it runs on."""

        def __init__(self, own_streams):
            self.scores = torch.empty(lib.pf_b200_ace_scores_bytes(w, h), dtype=torch.uint8, device=dev)
            self.minmax = torch.empty(lib.pf_b200_ace_minmax_bytes(w, h) // 8, dtype=torch.float64, device=dev)
            self.scratch = torch.empty(lib.pf_b200_canny_scratch_bytes(w, h), dtype=torch.uint8, device=dev)
            self.ace = torch.cuda.Stream(device=dev) if own_streams else None
            self.chain = torch.cuda.Stream(device=dev) if own_streams else None
            self.done = [torch.cuda.Event(), torch.cuda.Event()]

    slot0 = Slot(False)                  # everything on the current stream (stage timings)
    scores, minmax, scratch = slot0.scores, slot0.minmax, slot0.scratch

    def do_ace(i, sl=slot0, on=None):
        on = st if on is None else on
        args = (pages[i].data_ptr(), w, h, 0, h, xy.data_ptr(), 100, 10.0, 1000.0,
                sl.scores.data_ptr(), sl.minmax.data_ptr())
        chk(lib.pf_b200_dev_ace_minmax_init(sl.minmax.data_ptr(), w, h, on))
        chk(lib.pf_b200_dev_ace_scores(*args, on))
        chk(lib.pf_b200_dev_ace_refine(*args, on))
        chk(lib.pf_b200_dev_ace_scale(*args, out_ace[i].data_ptr(), on))

    def do_gauss(i, sl=slot0, on=None):
        chk(lib.pf_b200_dev_gaussian(pages[i].data_ptr(), out_g[i].data_ptr(), w, h, 2.0, 5,
                                     st if on is None else on))

    def do_sobel(i, sl=slot0, on=None):
        chk(lib.pf_b200_dev_sobel(out_g[i].data_ptr(), out_s[i].data_ptr(), w, h,
                                  st if on is None else on))

    def do_canny(i, sl=slot0, on=None):
        chk(lib.pf_b200_dev_canny(pages[i].data_ptr(), out_c[i].data_ptr(), w, h, 1,
                                  sl.scratch.data_ptr(), sl.scratch.numel(), st if on is None else on))

    stage_fns = [do_ace, do_gauss, do_sobel, do_canny]

    # Pages are independent, and so are ACE and the stencil chain of one page.
    # A step's ACE and its chain go on two streams, and consecutive steps
    # alternate between two such pairs (each with its own work areas), so the
    # latency-bound passes of one page (ACE's exact passes, Canny's hysteresis)
    # run while the next page's kernels fill the GPU.  Everything forks from the
    # current stream at the start of the timed region and joins it at the end:
    # the events that time the region are recorded on that stream.
    N_SLOTS = int(os.environ.get("PF_BENCH_SLOTS", "2"))
    slots = [Slot(True) for _ in range(N_SLOTS)]

    # (the stage breakdown below)
    def step_one_stream(i, marks=None):
        s = i % N_SETS
        for k, fn in enumerate(stage_fns):
            if marks is not None:
                marks[k][0].record()
            fn(s)
            if marks is not None:
                marks[k][1].record()

    def fork_all():
        main = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(main)
        for sl in slots:
            sl.ace.wait_event(ev)
            sl.chain.wait_event(ev)

    def join_all():
        main = torch.cuda.current_stream()
        for sl in slots:
            for k, stream in enumerate((sl.ace, sl.chain)):
                sl.done[k].record(stream)
                main.wait_event(sl.done[k])

    def step(i):
        s = i % N_SETS
        sl = slots[i % N_SLOTS]
        do_ace(s, sl, sl.ace.cuda_stream)
        do_gauss(s, sl, sl.chain.cuda_stream)
        do_sobel(s, sl, sl.chain.cuda_stream)
        do_canny(s, sl, sl.chain.cuda_stream)

    # ---- main timed region ----------------------------------------------------
    fork_all()
    for i in range(args.warmup):
        step(i)
    join_all()
    torch.cuda.synchronize()
    sampler.start()
    K = args.steps
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.synchronize()
    e0.record()
    fork_all()
    for i in range(K):
        step(args.warmup + i)
    join_all()
    e1.record()
    torch.cuda.synchronize()
    barrier()
    ms_step = e0.elapsed_time(e1) / K
    t = torch.tensor([ms_step], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item())
    value = world * npix / (ms_step * 1e-3) / 1e6

    # the same step with everything on one stream, stage by stage (where the
    # time goes; not the headline)
    K1 = max(5, min(K, 20))
    marks = [[[torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)]
              for _ in range(4)] for _ in range(K1)]
    for i in range(3):
        step_one_stream(i)
    torch.cuda.synchronize()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    for i in range(K1):
        step_one_stream(3 + i, marks[i])
    e3.record()
    torch.cuda.synchronize()
    ms_step_one = e2.elapsed_time(e3) / K1
    stage_ms = [float(np.mean([marks[i][k][0].elapsed_time(marks[i][k][1]) for i in range(K1)]))
                for k in range(4)]

    # ---- dominant kernel alone (ACE accumulation), for the roofline ----------
    def time_fn(fn, iters, warm=3):
        for i in range(warm):
            fn(i % N_SETS)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for i in range(iters):
            fn((warm + i) % N_SETS)
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / iters

    # (one page throughout: the three passes need each other's results)
    def ace_scores_only(i):
        chk(lib.pf_b200_dev_ace_minmax_init(minmax.data_ptr(), w, h, st))
        chk(lib.pf_b200_dev_ace_scores(pages[0].data_ptr(), w, h, 0, h, xy.data_ptr(), 100,
                                       10.0, 1000.0, scores.data_ptr(), minmax.data_ptr(), st))

    def ace_refine_only(i):
        chk(lib.pf_b200_dev_ace_refine(pages[0].data_ptr(), w, h, 0, h, xy.data_ptr(), 100,
                                       10.0, 1000.0, scores.data_ptr(), minmax.data_ptr(), st))

    def ace_scale_only(i):      # rescale + fix-up pass
        chk(lib.pf_b200_dev_ace_scale(pages[0].data_ptr(), w, h, 0, h, xy.data_ptr(), 100,
                                      10.0, 1000.0, scores.data_ptr(), minmax.data_ptr(),
                                      out_ace[0].data_ptr(), st))

    iters = max(10, min(K, 50))
    chk(lib.pf_b200_dev_ace_minmax_init(minmax.data_ptr(), w, h, st))
    ms_init = time_fn(lambda i: chk(lib.pf_b200_dev_ace_minmax_init(minmax.data_ptr(), w, h, st)), iters)
    ms_ace_scores = time_fn(ace_scores_only, iters) - ms_init
    ms_ace_refine = time_fn(ace_refine_only, iters)
    ms_ace_scale = time_fn(ace_scale_only, iters)
    # stencil stages on cold inputs: each iteration reads a page set last
    # touched N_SETS iterations (> 1 GB of traffic) ago
    ms_cold = [None, time_fn(do_gauss, iters), time_fn(do_sobel, iters), time_fn(do_canny, iters)]

    fp64_peak = lib.pf_b200_fp64_peak_tflops()
    fp32_peak = lib.pf_b200_fp32_peak_tflops()
    peaks = load_json(os.path.join(ROOT, "MEASURED_PEAKS.json"))
    hbm_peak = float(peaks["hbm_gbs"]) if peaks and "hbm_gbs" in peaks else 6650.0
    hbm_src = "MEASURED_PEAKS.json hbm_gbs" if peaks and "hbm_gbs" in peaks else "fallback 6650 GB/s"
    traffic = load_json(os.path.join(ROOT, "profiles", "traffic.json")) or {}
    ncu_row = (load_json(os.path.join(ROOT, "profiles", "r02_ncu_rows.json")) or {}).get("ace_scores_kernel", {})

    kept, skipped = count_pairs(w, h, xy_host)
    # What the kernel executes per pixel-sample pair that counts: d^2 (1 add), the
    # sum of weights (1 add), three saturating FMAs (the clamp) and three
    # accumulating FMAs: 8 instructions of the FP32 FMA pipe, 14 flops.
    ace_flops = kept * 14.0
    ace_tf = ace_flops / (ms_ace_scores * 1e-3) / 1e12
    ref_flops = kept * FLOPS_PAIR + skipped * FLOPS_SKIPPED
    roofline = {"kernel": "ace_scores_kernel", "bound": "fp32", "achieved": ace_tf,
                "peak": fp32_peak, "unit": "TFLOP/s", "frac": ace_tf / fp32_peak if fp32_peak > 0 else None,
                "traffic": traffic.get("ace_scores_kernel"),
                "peak_source": "pf_b200_fp32_peak_tflops(): FFMA probe run in this process "
                               "(MEASURED_PEAKS.json has no FP32 or FP64 figure)",
                "algorithmic_flops_per_launch": ace_flops, "kernel_ms": ms_ace_scores,
                "pairs_not_skipped": kept, "pairs_skipped": skipped,
                # FMA-pipe instruction slots (an add takes a slot like an FMA): 8 of them per pair
                "fma_pipe_slot_frac": (kept * 8.0 * 2.0 / (ms_ace_scores * 1e-3) / 1e12) / fp32_peak
                if fp32_peak > 0 else None,
                # one MUFU.RSQ per pair at the 16 lanes per clock and SM the XU pipe issues
                # (profiles/r02_pipe_rates.txt), at the SM clock the bench saw: the loop asks the
                # XU pipe, the FMA pipe and the issue port for about 8 cycles per pair each
                "mufu_floor_ms": None, "frac_of_mufu_floor": None,
                "issue_frac_ncu": ncu_row.get("issue_active_pct"),
                "fma_pipe_frac_ncu": ncu_row.get("pipe_fma_cycles_active_pct"),
                "xu_pipe_frac_ncu": ncu_row.get("pipe_xu_pct"),
                # the REFERENCE's arithmetic (22 FP64 operations per pair) retired per
                # second, against the FP64 FMA rate: round 1's figure of merit, kept for comparison
                "reference_fp64_flops_per_launch": ref_flops,
                "reference_fp64_rate_tflops": ref_flops / (ms_ace_scores * 1e-3) / 1e12,
                "fp64_peak_tflops": fp64_peak,
                "reference_fp64_rate_vs_fp64_peak": (ref_flops / (ms_ace_scores * 1e-3) / 1e12) / fp64_peak
                if fp64_peak > 0 else None,
                "note": "the accumulation pass runs in FP32 with a proven error bound and the pixels the bound "
                        "cannot decide are recomputed exactly (refine / fix-up passes, timed under "
                        "ace_kernels_ms): the result is bit-identical to the FP64 reference"}
    stencils = {}
    for k in (Stage.GAUSS, Stage.SOBEL, Stage.CANNY):
        gbs = 8.0 * npix / (ms_cold[k] * 1e-3) / 1e9
        stencils[Stage.NAMES[k]] = {
            "bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak,
            "traffic": traffic.get(Stage.NAMES[k]), "ms": ms_cold[k],
            "mpx_s": npix / (ms_cold[k] * 1e-3) / 1e6, "algorithmic_bytes_per_launch": 8 * npix}

    # ---- end to end through the host-pointer entry points ---------------------
    nbytes = npix * 4
    N_BLUR = 3                           # blurred pages in flight between the Gaussian and the Sobel thread
    pin = lib.pf_b200_host_alloc((4 + N_BLUR) * nbytes)
    if not pin:
        raise SystemExit("pinned allocation failed: " + _capi.last_error()[1])

    def host_view(k):
        return np.ctypeslib.as_array((ctypes.c_uint8 * nbytes).from_address(pin + k * nbytes)
                                     ).reshape(h, w, 4)

    hv = [host_view(k) for k in range(4 + N_BLUR)]
    hv[0][:] = host_pages[0]

    def bm(k):
        b = _capi.PfBitmap()
        b.size.x, b.size.y, b.pixels = w, h, pin + k * nbytes
        return b

    b_in, b_ace, b_g, b_s, b_c = [bm(k) for k in range(5)]
    b_blur = [b_g] + [bm(5 + k) for k in range(N_BLUR - 1)]
    xy_p = xy_host.ctypes.data_as(ctypes.POINTER(ctypes.c_int))

    def e2e_step():
        chk(lib.pf_b200_ace(ctypes.byref(b_in), ctypes.byref(b_ace), 100, xy_p, 10.0, 1000.0))
        chk(lib.pf_b200_gaussian(ctypes.byref(b_in), ctypes.byref(b_g), 2.0, 5))
        chk(lib.pf_b200_sobel(ctypes.byref(b_g), ctypes.byref(b_s)))
        chk(lib.pf_b200_canny(ctypes.byref(b_in), ctypes.byref(b_c)))

    # The same four calls from several host threads, as a caller with more than
    # one thread makes them (each thread has its own context in the library; a
    # call that runs beside others moves its page in one DMA each way, one that
    # is alone in row bands: pf_capi.cu, CallScope).
    #   headline e2e: one thread per call, each working through the pages in
    #     order with no barrier between steps -- the Sobel thread takes each
    #     blurred page from the Gaussian thread through a queue (N_BLUR host
    #     buffers), so every page still gets all four results, sobel(gaussian)
    #     included;
    #   beside it: two threads with a barrier after every step (round 2's first
    #     headline), and one thread.
    import queue

    def e2e_pages_in_flight(warm, steps):
        """Seconds for `steps` pages.  The threads are the same for the warm-up and the
        timed pages (a host thread's context in the library lives as long as the thread)."""
        ready, free, errs = queue.Queue(), threading.Semaphore(N_BLUR - 1), []
        gate = threading.Barrier(5)

        def each(call):
            def loop(n):
                for i in range(n):
                    call(i)
            return loop

        def gauss_loop(n):
            for i in range(n):
                free.acquire()
                chk(lib.pf_b200_gaussian(ctypes.byref(b_in), ctypes.byref(b_blur[i % N_BLUR]), 2.0, 5))
                ready.put(i)

        def sobel_loop(n):
            for _ in range(n):
                i = ready.get()
                chk(lib.pf_b200_sobel(ctypes.byref(b_blur[i % N_BLUR]), ctypes.byref(b_s)))
                free.release()

        def guarded(fn):
            def run():
                try:
                    torch.cuda.set_device(local)
                    fn(warm)
                    gate.wait()               # everyone has finished warming up
                    fn(steps)
                    gate.wait()               # everyone has finished the timed pages
                except Exception as exc:      # noqa: BLE001
                    errs.append(exc)
                    gate.abort()
                    ready.put(0)              # (do not leave the Sobel thread waiting)
                    free.release()
            return run

        loops = [each(lambda i: chk(lib.pf_b200_ace(ctypes.byref(b_in), ctypes.byref(b_ace), 100, xy_p,
                                                     10.0, 1000.0))),
                 gauss_loop, sobel_loop,
                 each(lambda i: chk(lib.pf_b200_canny(ctypes.byref(b_in), ctypes.byref(b_c))))]
        threads = [threading.Thread(target=guarded(fn), daemon=True) for fn in loops]
        for th in threads:
            th.start()
        t = None
        try:
            gate.wait()
            t0 = time.perf_counter()
            gate.wait()
            t = time.perf_counter() - t0
        except threading.BrokenBarrierError:
            pass
        for th in threads:
            th.join()
        if errs:
            raise errs[0]
        return t

    class HostWorker:
        """A host thread that makes the given calls, in order, once per step."""

        def __init__(self, calls):
            self.calls, self.todo, self.done = calls, queue.Queue(), queue.Queue()
            self.thread = threading.Thread(target=self.loop, daemon=True)
            self.thread.start()

        def loop(self):
            torch.cuda.set_device(local)
            while self.todo.get():
                try:
                    for call in self.calls:
                        call()
                    self.done.put(None)
                except Exception as exc:      # noqa: BLE001
                    self.done.put(exc)

        def stop(self):
            self.todo.put(False)
            self.thread.join()

    call_ace = lambda: chk(lib.pf_b200_ace(ctypes.byref(b_in), ctypes.byref(b_ace), 100, xy_p, 10.0, 1000.0))
    call_g = lambda: chk(lib.pf_b200_gaussian(ctypes.byref(b_in), ctypes.byref(b_g), 2.0, 5))
    call_s = lambda: chk(lib.pf_b200_sobel(ctypes.byref(b_g), ctypes.byref(b_s)))
    call_c = lambda: chk(lib.pf_b200_canny(ctypes.byref(b_in), ctypes.byref(b_c)))
    workers = [HostWorker([call_g, call_s])]

    def e2e_step_threads():
        for wk in workers:
            wk.todo.put(True)
        call_ace()
        call_c()
        for wk in workers:
            exc = wk.done.get()
            if exc is not None:
                raise exc

    def max_over_ranks(t):
        tt = torch.tensor([t], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())

    def timed_host_loop(fn, reps):
        for _ in range(max(3, args.warmup)):
            fn()
        barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        t = (time.perf_counter() - t0) / reps
        barrier()
        return max_over_ranks(t)

    # the extension that uploads the page once and returns all four results
    def e2e_chain_step():
        chk(lib.pf_b200_page_step(ctypes.byref(b_in), ctypes.byref(b_ace), ctypes.byref(b_g), ctypes.byref(b_s),
                                  ctypes.byref(b_c), 100, xy_p, 10.0, 1000.0, 2.0, 5))

    ke = max(5, min(K, 30))
    t_e2e_chain = timed_host_loop(e2e_chain_step, ke)
    chain_views = [hv[k + 1].copy() for k in range(4)] if rank == 0 else None
    t_e2e_one = timed_host_loop(e2e_step, ke)
    t_e2e_two = timed_host_loop(e2e_step_threads, ke)
    for wk in workers:
        wk.stop()
    # (three runs of ke pages, the median reported: 30 pages are 0.1 s of wall clock,
    # and host threads are at the mercy of the box's scheduler)
    e2e_runs = []
    for _ in range(3):
        barrier()
        t_run = e2e_pages_in_flight(max(3, args.warmup), ke) / ke
        barrier()
        e2e_runs.append(max_over_ranks(t_run))
    t_e2e = sorted(e2e_runs)[1]
    blur_copies_equal = all(bool(np.array_equal(hv[2], hv[5 + k])) for k in range(N_BLUR - 1))
    e2e_value = world * npix / t_e2e / 1e6
    torch.cuda.synchronize()
    clocks = sampler.stop()
    sm_mhz = (clocks or {}).get("sm_mhz") or 0
    if sm_mhz:
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        roofline["mufu_floor_ms"] = kept / (16.0 * sms * sm_mhz * 1e6) * 1e3
        roofline["frac_of_mufu_floor"] = roofline["mufu_floor_ms"] / ms_ace_scores

    # the same four calls on ordinary (malloc'ed, pageable) bitmaps on both sides:
    # what an existing C program passes
    pg = [np.empty((h, w, 4), dtype=np.uint8) for _ in range(5)]
    pg[0][:] = host_pages[0]

    def bmp(a):
        b = _capi.PfBitmap()
        b.size.x, b.size.y, b.pixels = w, h, a.ctypes.data
        return b

    p_in, p_ace, p_g, p_s, p_c = [bmp(a) for a in pg]

    def e2e_pageable_step():
        chk(lib.pf_b200_ace(ctypes.byref(p_in), ctypes.byref(p_ace), 100, xy_p, 10.0, 1000.0))
        chk(lib.pf_b200_gaussian(ctypes.byref(p_in), ctypes.byref(p_g), 2.0, 5))
        chk(lib.pf_b200_sobel(ctypes.byref(p_g), ctypes.byref(p_s)))
        chk(lib.pf_b200_canny(ctypes.byref(p_in), ctypes.byref(p_c)))

    for _ in range(3):
        e2e_pageable_step()
    kp = max(5, min(K, 20))
    barrier()
    t0 = time.perf_counter()
    for _ in range(kp):
        e2e_pageable_step()
    t_pageable = (time.perf_counter() - t0) / kp
    barrier()
    tp = torch.tensor([t_pageable], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tp, op=dist.ReduceOp.MAX)
    t_pageable = float(tp.item())

    # All four e2e results (both paths) against the CPU oracle, outside the timed
    # regions.  ACE at full size costs the oracle ~10 s on 16 cores: rank 0 only.
    e2e_ok = None
    if rank == 0 and not args.no_cpu_baseline:
        from oracle import pf_oracle as o
        cores = os.cpu_count() or 1
        ref_g = o.gaussian(host_pages[0], 2.0, 5)
        refs = [o.ace_with_samples(host_pages[0], xy_host, 10.0, 1000.0, nb_threads=cores),
                ref_g, o.sobel(ref_g), o.canny(host_pages[0])]
        e2e_ok = {"pinned": [bool(np.array_equal(hv[k + 1], refs[k])) for k in range(4)],
                  "pageable": [bool(np.array_equal(pg[k + 1], refs[k])) for k in range(4)],
                  "order": ["ace", "gaussian", "sobel", "canny"],
                  "against": "oracle/pf_oracle.c (bit for bit)"}
        e2e_ok["page_step"] = [bool(np.array_equal(chain_views[k], refs[k])) for k in range(4)]
        e2e_ok["blurred_pages_in_flight_equal"] = blur_copies_equal
        e2e_ok["all"] = (all(e2e_ok["pinned"]) and all(e2e_ok["pageable"]) and all(e2e_ok["page_step"])
                         and blur_copies_equal)

    # ---- extras: ACE 600 dpi, 400 samples, row bands over the ranks -----------
    extras = {}
    if not args.no_extras:
        w6, h6 = A4_600
        big = torch.empty((h6, w6, 4), dtype=torch.uint8, device=dev)
        if rank == 0:
            small = torch.from_numpy(host_pages[0]).to(dev)
            big[:] = small.repeat_interleave(2, dim=0).repeat_interleave(2, dim=1)[:h6, :w6]
        xy6_host = pf.ace_samples(400, w6, h6, ACE_SEED)
        xy6 = xy_tensor(xy6_host, dev)
        comp = CudaAceCompute()
        for _ in range(2):
            ace_sharded(big, xy6, 10.0, 1000.0, comp)
        n6 = 5
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        torch.cuda.synchronize()
        a.record()
        for _ in range(n6):
            out6 = ace_sharded(big, xy6, 10.0, 1000.0, comp)
        b.record()
        torch.cuda.synchronize()
        barrier()
        t6 = torch.tensor([a.elapsed_time(b) / n6], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t6, op=dist.ReduceOp.MAX)
        # the split must not change the result: rank 0 (which holds the page)
        # redoes it alone and compares bit for bit
        same6 = None
        if rank == 0:
            mm6 = comp.new_minmax(dev, w6, h6)
            sc6 = comp.scores(big, xy6, 10.0, 1000.0, (0, h6), mm6)
            comp.refine(big, xy6, 10.0, 1000.0, (0, h6), sc6, mm6)
            same6 = bool(torch.equal(comp.scale(big, xy6, 10.0, 1000.0, (0, h6), sc6, mm6), out6))
            del sc6
        extras["ace_600dpi_row_bands"] = {
            "workload": "pf_apply_ace 4960x7016, 400 samples, rows split over %d GPU(s); NCCL scatter "
                        "of the bands + broadcast of the samples' pixel values + one exchange of the "
                        "exact extremes (6 + 6 doubles) + band all-gather inside the timed region" % world,
            "ms": float(t6.item()), "mpx_s": w6 * h6 / (float(t6.item()) * 1e-3) / 1e6,
            "scaling": "strong", "equals_single_gpu_result": same6}
        # the same over NVLink peer memory: no scatter, no all-gather
        if world > 1:
            try:
                from libpillowfight_b200.distributed import AcePeerBands
                peer = AcePeerBands(h6, w6, dev)
                if rank == 0:
                    peer.page.copy_(big)
                for _ in range(2):
                    outp = peer.run(xy6, 10.0, 1000.0)
                barrier()
                torch.cuda.synchronize()
                a.record()
                for _ in range(n6):
                    outp = peer.run(xy6, 10.0, 1000.0)
                b.record()
                torch.cuda.synchronize()
                barrier()
                tp6 = torch.tensor([a.elapsed_time(b) / n6], dtype=torch.float64, device=dev)
                dist.all_reduce(tp6, op=dist.ReduceOp.MAX)
                okp = torch.tensor([1 if torch.equal(outp, out6) else 0], device=dev)
                dist.all_reduce(okp, op=dist.ReduceOp.MIN)
                extras["ace_600dpi_row_bands_peer"] = {
                    "workload": "the same page; every GPU reads its band out of rank 0's page and the "
                                "rescale kernel writes it into every GPU's result over NVLink peer "
                                "memory (torch symmetric memory); NCCL only for the sample values and "
                                "the one exchange of the extremes",
                    "ms": float(tp6.item()), "mpx_s": w6 * h6 / (float(tp6.item()) * 1e-3) / 1e6,
                    "scaling": "strong", "equals_nccl_result_on_every_rank": bool(okp.item()),
                    "result_stores": "multimem.st to an NVLS multicast address" if peer.use_multicast
                                     else "one st.global per peer"}
                if peer.use_multicast:      # and with plain peer stores, for comparison
                    peer.use_multicast = False
                    for _ in range(2):
                        outp = peer.run(xy6, 10.0, 1000.0)
                    barrier()
                    torch.cuda.synchronize()
                    a.record()
                    for _ in range(n6):
                        outp = peer.run(xy6, 10.0, 1000.0)
                    b.record()
                    torch.cuda.synchronize()
                    barrier()
                    tq6 = torch.tensor([a.elapsed_time(b) / n6], dtype=torch.float64, device=dev)
                    dist.all_reduce(tq6, op=dist.ReduceOp.MAX)
                    okq = torch.tensor([1 if torch.equal(outp, out6) else 0], device=dev)
                    dist.all_reduce(okq, op=dist.ReduceOp.MIN)
                    extras["ace_600dpi_row_bands_peer"]["ms_with_per_peer_stores"] = float(tq6.item())
                    extras["ace_600dpi_row_bands_peer"]["per_peer_stores_equal"] = bool(okq.item())
                del peer, outp
            except Exception as exc:      # no peer access / no symmetric memory on this box
                extras["ace_600dpi_row_bands_peer"] = {"unavailable": repr(exc)[:300]}
        del big, out6

        # ---- extras: page batch through the stencil chain (BASELINE.json config 5:
        # 512 pages over 8 GPUs = 64 pages per GPU; pages are independent, no collective)
        nbp = 64
        batch = torch.empty((nbp, h, w, 4), dtype=torch.uint8, device=dev)
        for i in range(nbp):
            batch[i] = pages[i % N_SETS]
        bg, bs, bc = torch.empty_like(batch), torch.empty_like(batch), torch.empty_like(batch)
        chain = CudaChain()

        def chain_pass():
            chain.gaussian(batch, bg)
            chain.sobel(bg, bs)
            chain.canny(batch, bc, pages_per_launch=8)

        chain_pass()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        torch.cuda.synchronize()
        a.record()
        for _ in range(3):
            chain_pass()
        b.record()
        torch.cuda.synchronize()
        barrier()
        tb = torch.tensor([a.elapsed_time(b) / 3], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tb, op=dist.ReduceOp.MAX)
        extras["chain_batch"] = {
            "workload": "%d A4 pages per GPU (%d in all) through pf_b200_dev_gaussian_batch -> "
                        "sobel_batch, and pf_b200_dev_canny on the pages; device-resident, "
                        "pages sharded over the GPUs, no collective" % (nbp, nbp * world),
            "ms": float(tb.item()),
            "mpx_s": world * nbp * npix / (float(tb.item()) * 1e-3) / 1e6,
            "hbm_gbs_of_24B_per_px": world * nbp * npix * 24 / (float(tb.item()) * 1e-3) / 1e9,
            "scaling": "weak"}
        del batch, bg, bs, bc

    # ---- CPU baseline (rank 0, N = 1 only) ------------------------------------
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import pf_oracle as o
        o.lib()
        cores = os.cpu_count() or 1
        npg = cpu_pages_per_step(cores)
        cpu_pages = [host_pages[i % 2] for i in range(npg)]
        rows = cpu_sample_rows(o, cpu_pages, xy_host, cores, 1, 20.0)
        secs = cpu_step(o, cpu_pages, xy_host, cores, rows)
        cpu_baseline = {
            "value": npg * w * min(rows, h) / secs / 1e6, "unit": "Mpx/s", "cores": cores,
            "kind": "port",
            "sample": "one step of %d A4 300-dpi page(s), top %d of %d rows each, %.1f s: oracle ACE "
                      "with %d pthreads page after page, then Gaussian->Sobel and Canny one page "
                      "per thread (the reference's stencils are single-threaded)"
                      % (npg, min(rows, h), h, secs, cores)}

    lib.pf_b200_host_free(pin)
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "Mpx/s",
            "n_gpus": world, "steps": K, "warmup": args.warmup, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32+f64", "data": "synthetic",
            "config": dict(CONFIG),
            "details": {
                "pages_per_step": world,
                "streams": "a step's ACE on one stream, its Gaussian->Sobel and Canny on a second; consecutive "
                           "steps alternate between two such pairs of streams (own work areas each); all "
                           "fork from and join the timing stream at the ends of the timed region",
                "ms_per_step_one_stream": ms_step_one,
                "l2": "inputs rotate over %d page sets per GPU (%.1f GB in+out), so a page is not "
                      "L2-resident when its step starts" % (N_SETS, N_SETS * 5 * nbytes / 1e9),
                "hbm_peak_source": hbm_src,
                "dtype": "ACE: FP32 accumulation with a proven error bound, FP64 and the reference's own "
                         "FP64 operation order for the pixels the bound leaves open (bit-identical to "
                         "the FP64 reference); Gaussian: FP64; Sobel, Canny: integer"},
            "roofline": roofline,
            "roofline_stencils": stencils,
            "stages": {Stage.NAMES[k]: {"ms_in_step": stage_ms[k],
                                        "mpx_s": npix / (stage_ms[k] * 1e-3) / 1e6,
                                        "share": stage_ms[k] / sum(stage_ms)} for k in range(4)},
            "ace_kernels_ms": {"scores": ms_ace_scores, "refine": ms_ace_refine,
                               "scale_and_fixup": ms_ace_scale,
                               "pixels_recomputed_per_page": int(scores[:256].view(torch.int32)[32])},
            "cpu_baseline": cpu_baseline,
            "e2e": {"value": e2e_value, "unit": "Mpx/s", "h2d_bytes_per_step": 4 * nbytes + xy_host.nbytes,
                    "d2h_bytes_per_step": 4 * nbytes, "ms_per_step": t_e2e * 1e3,
                    "api": "pf_b200_ace/gaussian/sobel/canny on pinned host bitmaps, one host thread per "
                           "call, each working through the pages in order (pages in flight; the Sobel "
                           "thread takes every blurred page from the Gaussian thread, so each page gets "
                           "all four results of the step)",
                    "host_threads": 4, "runs_ms_per_step": [t * 1e3 for t in e2e_runs],
                    "two_threads_step_at_a_time": {"value": world * npix / t_e2e_two / 1e6, "unit": "Mpx/s",
                                                   "ms_per_step": t_e2e_two * 1e3},
                    "one_thread": {"value": world * npix / t_e2e_one / 1e6, "unit": "Mpx/s",
                                   "ms_per_step": t_e2e_one * 1e3},
                    "result_checked": e2e_ok},
            "e2e_page_step": {"value": world * npix / t_e2e_chain / 1e6, "unit": "Mpx/s",
                              "ms_per_step": t_e2e_chain * 1e3, "h2d_bytes_per_step": nbytes + xy_host.nbytes,
                              "d2h_bytes_per_step": 4 * nbytes,
                              "api": "pf_b200_page_step (an extension, not the drop-in interface): one upload, "
                                     "the four results of the step"},
            "e2e_pageable": {"value": world * npix / t_pageable / 1e6, "unit": "Mpx/s",
                             "ms_per_step": t_pageable * 1e3,
                             "api": "the same four calls on malloc'ed (pageable) bitmaps, in and out"},
            "gpu_launches": K * sum(Stage.LAUNCHES),
            "clocks": clocks,
            "extras": extras,
        }
        emit(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


_REAL_STDOUT = None


def emit(line):
    """The one JSON line, on the process's real stdout."""
    data = (line + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(line + "\n")
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    args = parse_args()
    # Libraries (NCCL's version banner, for one) print to stdout; the contract is
    # ONE JSON line there.  Everything else is sent to stderr.
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
