"""Both directions at once, one page each way, in chunks: does the chunk size or an
event after every chunk cost bandwidth?"""
import torch
n = 2480 * 3508 * 4
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s_up, s_down = torch.cuda.Stream(), torch.cuda.Stream()

def run(chunks, events, up=True, down=True, reps=10):
    c = n // chunks
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); s_up.wait_event(a); s_down.wait_event(a)
    for _ in range(reps):
        for k in range(chunks):
            if up:
                with torch.cuda.stream(s_up):
                    d_in[k * c:(k + 1) * c].copy_(h_in[k * c:(k + 1) * c], non_blocking=True)
                    if events:
                        torch.cuda.Event().record(s_up)
            if down:
                with torch.cuda.stream(s_down):
                    h_out[k * c:(k + 1) * c].copy_(d_out[k * c:(k + 1) * c], non_blocking=True)
                    if events:
                        torch.cuda.Event().record(s_down)
    e1, e2 = torch.cuda.Event(), torch.cuda.Event()
    e1.record(s_up); e2.record(s_down)
    torch.cuda.current_stream().wait_event(e1); torch.cuda.current_stream().wait_event(e2)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps

for chunks in (1, 2, 4, 8, 16):
    for events in (False, True):
        run(chunks, events)
        t = run(chunks, events)
        tu = run(chunks, events, down=False)
        print("%2d chunks%s: both ways %.3f ms (%.1f GB/s each), up alone %.3f ms" % (
            chunks, ", event after each" if events else "", t, n / t / 1e6, tu), flush=True)
