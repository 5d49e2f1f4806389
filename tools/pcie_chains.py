"""S streams, each a chain of [page up -> small kernel -> page down], queued deep
with no host synchronisation: what dependent DMAs reach without host round trips."""
import sys, torch
n = 2480 * 3508 * 4
K = 20
for S in (1, 2, 4, 8):
    hi = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(S)]
    ho = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(S)]
    di = [torch.empty(n, dtype=torch.uint8, device="cuda") for _ in range(S)]
    do = [torch.empty(n, dtype=torch.uint8, device="cuda") for _ in range(S)]
    st = [torch.cuda.Stream() for _ in range(S)]
    for split in (False, True):
        # split: upload and download on separate streams joined by events (as the library does)
        st2 = [torch.cuda.Stream() for _ in range(S)]
        def go(reps):
            for _ in range(reps):
                for s in range(S):
                    with torch.cuda.stream(st[s]):
                        di[s].copy_(hi[s], non_blocking=True)
                        do[s][:4096].copy_(di[s][:4096])
                        if split:
                            e = torch.cuda.Event(); e.record(st[s]); st2[s].wait_event(e)
                    with torch.cuda.stream(st2[s] if split else st[s]):
                        ho[s].copy_(do[s], non_blocking=True)
        go(2)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for s in st + st2: s.wait_event(a)
        go(K)
        for s in st + st2:
            e = torch.cuda.Event(); e.record(s); torch.cuda.current_stream().wait_event(e)
        b.record()
        torch.cuda.synchronize()
        t = a.elapsed_time(b) / (K * S)
        print("%d chains%s: %.3f ms per page pair (%.1f GB/s each way)" % (S, " (down on a second stream)" if split else "", t, n / t / 1e6), flush=True)
