#!/usr/bin/env python
"""Host<->device copy rates of this box (pinned memory, one copy stream each way), the denominator for
the e2e number of bench.py: the fused path uploads 252 MB and downloads 117 MB per step."""
import json
import torch

dev = torch.device("cuda:0")
n = 256 << 20
h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
d = torch.empty(n, dtype=torch.uint8, device=dev)
h2, d2 = torch.empty(n, dtype=torch.uint8, pin_memory=True), torch.empty(n, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def rate(fn, streams, reps=10):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for s in streams:
        s.wait_event(a)
    for _ in range(reps):
        fn()
    for s in streams:
        torch.cuda.current_stream().wait_stream(s)
    b.record()
    torch.cuda.synchronize()
    return n * reps / (a.elapsed_time(b) * 1e-3) / 1e9


def up():
    with torch.cuda.stream(s1):
        d.copy_(h, non_blocking=True)


def down():
    with torch.cuda.stream(s2):
        h2.copy_(d2, non_blocking=True)


def both():
    up()
    down()


print(json.dumps({"h2d_GBps": rate(up, [s1]), "d2h_GBps": rate(down, [s2]), "h2d_GBps_while_d2h": rate(both, [s1, s2]),
                  "bytes_per_copy": n}))
