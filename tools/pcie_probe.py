"""PCIe ceiling for the host-pointer calls: pinned copies of one A4 page, each
direction alone and both at once (the band pipeline overlaps them)."""
import torch

n = 2480 * 3508 * 4
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s_up, s_down = torch.cuda.Stream(), torch.cuda.Stream()


def run(up, down, reps=20):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    s_up.wait_event(a)
    s_down.wait_event(a)
    for _ in range(reps):
        if up:
            with torch.cuda.stream(s_up):
                d_in.copy_(h_in, non_blocking=True)
        if down:
            with torch.cuda.stream(s_down):
                h_out.copy_(d_out, non_blocking=True)
    e1, e2 = torch.cuda.Event(), torch.cuda.Event()
    e1.record(s_up)
    e2.record(s_down)
    torch.cuda.current_stream().wait_event(e1)
    torch.cuda.current_stream().wait_event(e2)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


for _ in range(2):
    t_up, t_down, t_both = run(True, False), run(False, True), run(True, True)
print("one A4 page (%.1f MB): up %.3f ms (%.1f GB/s), down %.3f ms (%.1f GB/s), both at once %.3f ms (%.1f GB/s each way)"
      % (n / 1e6, t_up, n / t_up / 1e6, t_down, n / t_down / 1e6, t_both, n / t_both / 1e6))
