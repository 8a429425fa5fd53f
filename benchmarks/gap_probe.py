#!/usr/bin/env python
"""Where does the step time outside the scoring kernels go?  Runs the bench.py step loop (B = 8 Qwen2-0.5B
windows, S = 4096, fused scoring) in several pipeline configurations and prints ms/step."""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "llm-inference-in-distributed-edge-networks_b200"))
import torch
import splitq
from splitq import _native as N

B, H, KVH, S, D = 8, 14, 2, 4096, 896
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
q = torch.randn(B, H, S, 64, device=dev, generator=g); k = torch.randn(B, KVH, S, 64, device=dev, generator=g)
h = torch.randn(B, S, D, device=dev, generator=g)
ladder = splitq.ladder_from_fractions([(0.25, 2), (0.25, 4), (0.25, 8)], S)
mk = lambda: splitq.BoundaryPlan(B, H, S, D, ladder, scale_mode=N.SCALE_TOKEN, KVH=KVH, src_device=dev)
STEPS = 40

def timed(fn, steps=STEPS, warm=5):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / steps

plan = mk()
print(f"score only, one stream              {timed(lambda: plan.score_fused(q, k)):.3f} ms")
print(f"whole path serial on one stream     {timed(lambda: plan.run_fused(q, k, h)):.3f} ms")
def side_only():
    plan.send(h); plan.hop(); plan.receive()
print(f"select+pack+hop+dequant only        {timed(side_only):.3f} ms")
for name, kw, ev in [("overlapped, prio -3, events", dict(main_priority=-3), True),
                     ("overlapped, prio -3, no events", dict(main_priority=-3), False),
                     ("overlapped, caller stream", dict(main_priority=None), False),
                     ("overlapped, prio -3, pack on main", dict(main_priority=-3, pack_on_side=False), False),
                     ("overlapped, prio -3, depth 3", dict(main_priority=-3, depth=3), False)]:
    ob = splitq.OverlappedBoundary(mk, fused=True, **({"depth": 2} | kw))
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(STEPS + 5)]
    it = iter(evs)
    def step():
        ob.step(q, k, h, score_events=next(it) if ev else None)
    for _ in range(5): step()
    ob.drain(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); a.record()
    for _ in range(STEPS): step()
    t1 = time.perf_counter(); ob.join(); b.record(); torch.cuda.synchronize()
    msg = f"{name:36s}{a.elapsed_time(b) / STEPS:.3f} ms/step  host submit {1e3 * (t1 - t0) / STEPS:.3f} ms/step"
    if ev:
        used = evs[5:5 + STEPS]
        sc = sum(x.elapsed_time(y) for x, y in used) / STEPS
        gap = sum(used[i - 1][1].elapsed_time(used[i][0]) for i in range(1, STEPS)) / (STEPS - 1)
        msg += f"  score {sc:.3f}  gap {gap:.3f}"
    print(msg)
