"""Pinned host <-> device copy rates of the box (what bounds bench.py's e2e leg): prints GB/s for D2H, H2D and both at once."""
import torch

n = 1 << 30
d = torch.empty(n, dtype=torch.uint8, device="cuda")
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=4):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return reps * n / (a.elapsed_time(b) * 1e-3) / 1e9


print("D2H GB/s", round(timed(lambda: h.copy_(d, non_blocking=True)), 1))
print("H2D GB/s", round(timed(lambda: d.copy_(h, non_blocking=True)), 1))


def both():
    e = torch.cuda.Event()
    e.record()
    s1.wait_event(e)
    s2.wait_event(e)
    with torch.cuda.stream(s1):
        h.copy_(d, non_blocking=True)
    with torch.cuda.stream(s2):
        d2.copy_(h2, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1)
    torch.cuda.current_stream().wait_stream(s2)


print("D2H+H2D together, GB/s each way", round(timed(both), 1))
