"""Developer probe: pinned host <-> device copy bandwidth of this box (the ceiling of bench.py's e2e number)."""
import torch
n = 1 << 30
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(n // 2, dtype=torch.uint8).pin_memory()
d2 = torch.empty(n // 2, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def t(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


print("H2D alone   %.1f GB/s" % (n / t(lambda: d.copy_(h, non_blocking=True)) / 1e6))
print("D2H alone   %.1f GB/s" % (n / t(lambda: h.copy_(d, non_blocking=True)) / 1e6))


def both():
    e = torch.cuda.Event()
    e.record()
    with torch.cuda.stream(s1):
        s1.wait_event(e)
        d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2):
        s2.wait_event(e)
        h2.copy_(d2, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1)
    torch.cuda.current_stream().wait_stream(s2)


ms = t(both)
print("H2D 1 GiB + D2H 0.5 GiB concurrently: %.2f ms -> H2D %.1f GB/s, D2H %.1f GB/s" % (ms, n / ms / 1e6, n / 2 / ms / 1e6))
