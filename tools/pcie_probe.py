"""PCIe ceiling of the box for the e2e byte mix: pinned D2H alone, H2D alone, and both directions at once
(two streams), 1 GiB per copy, CUDA events.  Explains the e2e number of bench.py (D2H-bound)."""
import torch

dev = torch.device("cuda:0")
n = 1 << 30
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_a = torch.empty(n, dtype=torch.uint8, device=dev)
d_b = torch.empty(n, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    for s in (s1, s2):
        torch.cuda.current_stream().wait_stream(s)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def d2h():
    with torch.cuda.stream(s1):
        h_out.copy_(d_a, non_blocking=True)


def h2d():
    with torch.cuda.stream(s2):
        d_b.copy_(h_in, non_blocking=True)


def both():
    d2h(); h2d()


for name, fn, nbytes in (("D2H alone", d2h, n), ("H2D alone", h2d, n), ("D2H + H2D concurrently (per direction)", both, n)):
    ms = timed(fn)
    print(f"{name:42s} {nbytes / ms / 1e6:7.1f} GB/s")
