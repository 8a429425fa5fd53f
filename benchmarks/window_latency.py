#!/usr/bin/env python
"""Per-window latency of the boundary path at the reference's batch size (B = 1 window), with
eager launches and replayed as one CUDA graph (SURVEY 7 hard part 4: at B = 1 the path is
launch-latency-bound, so latency is reported separately from the batched throughput of bench.py)."""
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "llm-inference-in-distributed-edge-networks_b200"))
import torch  # noqa: E402

import splitq  # noqa: E402
from splitq import _native as N  # noqa: E402

CASES = [("pythia-70m S=512 group int4", 8, 8, 512, 512, [(0.5, 4)], N.SCALE_GROUP),
         ("pythia-70m S=2048 token int4", 8, 8, 2048, 512, [(0.5, 4)], N.SCALE_TOKEN),
         ("qwen2-0.5b S=512 group int4", 14, 2, 512, 896, [(0.5, 4)], N.SCALE_GROUP),
         ("qwen2-0.5b S=4096 ladder 8/4/2", 14, 2, 4096, 896, [(0.25, 2), (0.25, 4), (0.25, 8)], N.SCALE_TOKEN)]


def timed(fn, iters=200):
    for _ in range(20):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return statistics.median(ts)


def main():
    dev = "cuda:0"
    gen = torch.Generator(device=dev).manual_seed(0)
    out = []
    for name, H, KVH, S, D, fracs, mode in CASES:
        ladder = splitq.ladder_from_fractions(fracs, S)
        plan = splitq.BoundaryPlan(1, H, S, D, ladder, scale_mode=mode, KVH=KVH)
        q = torch.randn(1, H, S, 64, device=dev, generator=gen)
        k = torch.randn(1, KVH, S, 64, device=dev, generator=gen)
        h = torch.randn(1, S, D, device=dev, generator=gen)
        want = plan.run_fused(q, k, h).hidden.clone()
        eager_us = timed(lambda: plan.run_fused(q, k, h))
        graph = plan.capture(q, k, h, fused=True)
        graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(plan.h_out, want), "graph replay differs from eager"
        graph_us = timed(graph.replay)
        rec = {"case": name, "eager_us_per_window": eager_us, "graph_us_per_window": graph_us,
               "launches": plan.launches_per_run(True)}
        print(json.dumps(rec), flush=True)
        out.append(rec)
    with open(os.path.join(ROOT, "profiles", "r1_window_latency.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
