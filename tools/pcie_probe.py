#!/usr/bin/env python
"""PCIe copy rates of the box for the end-to-end leg's transfer sizes (pinned host memory, copy engines):
D2H of one frame's images (1.2 + 3.7 + 3.7 MB as three copies and as one 8.6 MB copy), H2D of a depth image,
alone and with both directions at once."""
import json
import torch

dev = torch.device("cuda", 0)
n = 640 * 480
d_small = torch.zeros(n, dtype=torch.float32, device=dev)
d_big = [torch.zeros(3 * n, dtype=torch.float32, device=dev) for _ in range(2)]
d_all = torch.zeros(7 * n, dtype=torch.float32, device=dev)
h_small = torch.zeros(n, dtype=torch.float32).pin_memory()
h_big = [torch.zeros(3 * n, dtype=torch.float32).pin_memory() for _ in range(2)]
h_all = torch.zeros(7 * n, dtype=torch.float32).pin_memory()
h_up = torch.zeros(n, dtype=torch.float32).pin_memory()
d_up = torch.zeros(n, dtype=torch.float32, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(s1)
    for _ in range(reps):
        fn()
    e1.record(s1)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


def d2h3():
    with torch.cuda.stream(s1):
        h_small.copy_(d_small, non_blocking=True)
        h_big[0].copy_(d_big[0], non_blocking=True)
        h_big[1].copy_(d_big[1], non_blocking=True)


def d2h1():
    with torch.cuda.stream(s1):
        h_all.copy_(d_all, non_blocking=True)


def h2d():
    with torch.cuda.stream(s1):
        d_up.copy_(h_up, non_blocking=True)


def both():
    with torch.cuda.stream(s2):
        d_up.copy_(h_up, non_blocking=True)
    d2h3()


out = {}
for name, fn, nbytes in (("d2h_three_copies", d2h3, 28 * n), ("d2h_one_copy", d2h1, 28 * n), ("h2d_depth", h2d, 4 * n),
                         ("d2h_three_copies_with_h2d", both, 28 * n)):
    us = timed(fn)
    out[name] = dict(us=round(us, 1), gbs=round(nbytes / us / 1e3, 1))
print(json.dumps(out))
