#!/usr/bin/env python
"""Split-inference evaluation throughput (BASELINE.json: "eval tok/s 1-8 GPU") next to the
host-CPU port of the reference loop.

    python benchmarks/eval_throughput.py --family qwen [--windows 8]
    torchrun --nproc-per-node 2 benchmarks/eval_throughput.py --family qwen --windows 16

Per window the sweep is the notebook's: M methods x L split layers x R ratios (default 3 x 5 x 5 =
75 split forwards, Qwen2-0.5B, S=512, stride 32; Pythia-70M uses S=2048 and 4 layers like its
params.json).  Reported: window tokens/s (windows*S / time), split-forward token throughput
(x M*L*R; comparable to the reference notebook's ~3.2 k token-forwards/s on a T4), and the same
two figures for the CPU port on --cpu-windows windows (rank 0 only; run through
`bench.py --impl reference --eval-family ...`, the one place that executes oracle/).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "llm-inference-in-distributed-edge-networks_b200"))

import torch  # noqa: E402

from splitq import compat, evalloop  # noqa: E402
from splitq.models import Pythia70Model, QwenPointFiveBModel  # noqa: E402

PARAMS = {
    "qwen": dict(ratios=[0, 0.25, 0.5, 0.75, 1], layers_of_interest=[3, 11, 18, 22, 23],
                 methods=["regular_importance", "last_row", "aggregate_till"], max_length=512, stride=32),
    "pythia": dict(ratios=[0, 0.25, 0.5, 0.75, 1], layers_of_interest=[1, 2, 4, 5],
                   methods=["regular_importance", "last_row", "aggregate_till"], max_length=2048, stride=32),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--family", default="qwen", choices=["qwen", "pythia"])
    ap.add_argument("--windows", type=int, default=8, help="windows in the token stream (all ranks together)")
    ap.add_argument("--cpu-windows", type=int, default=1)
    ap.add_argument("--importance", default="fused", choices=["fused", "materialised"])
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--batch-methods", action="store_true", help="methods of one (layer, ratio) as one downstream batch")
    ap.add_argument("--batch-combos", action="store_true", help="all (method, ratio > 0) variants of a layer as one batch")
    args = ap.parse_args()
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=torch.device(dev))
    params = PARAMS[args.family]
    S, stride = params["max_length"], params["stride"]
    if args.family == "qwen":
        cfg = compat.QWEN2_05B
        wrapper = QwenPointFiveBModel.from_config(dev, cfg=cfg, seed=0)
    else:
        cfg = compat.PYTHIA_70M
        wrapper = Pythia70Model.from_config(dev, params["ratios"], cfg=cfg, seed=0)
    n_tokens = S + (args.windows - 1) * stride
    ids = torch.randint(0, cfg["vocab_size"], (1, n_tokens), generator=torch.Generator().manual_seed(1234))
    fn = evalloop.make_importance_fn(wrapper, params["methods"], args.importance, layers=params["layers_of_interest"])
    kw = dict(reduce_device=dev if world > 1 else None, log=lambda *a: None, batch_methods=args.batch_methods,
              batch_combos=args.batch_combos)
    evaluate = lambda: evaluate_once(params, wrapper, ids, fn, S, kw)  # noqa: E731
    evaluate()  # warm-up (cuBLAS heuristics, workspaces)
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    t0 = time.perf_counter()
    res = evaluate()
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    dt = time.perf_counter() - t0
    combos = len(params["methods"]) * len(params["layers_of_interest"]) * len(params["ratios"])
    out = {"family": args.family, "n_gpus": world, "windows": res.windows, "S": S, "combos_per_window": combos,
           "importance_impl": args.importance, "batch_methods": args.batch_methods, "batch_combos": args.batch_combos, "seconds": dt, "window_tokens_per_s": res.windows * S / dt,
           "split_forward_tokens_per_s": res.windows * S * combos / dt, "ppl_no_quant": res.ppl[0][0][0]}
    if rank == 0 and not args.no_cpu:
        # the CPU port of the reference loop lives in oracle/ and is only ever executed by bench.py
        import subprocess
        line = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--eval-family",
                               args.family, "--steps", str(args.cpu_windows)], capture_output=True, text=True,
                              check=True).stdout.strip().splitlines()[-1]
        cpu = json.loads(line)
        out["cpu_port"] = {k: cpu[k] for k in ("cores", "windows", "seconds", "window_tokens_per_s",
                                                "split_forward_tokens_per_s")}
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        torch.distributed.destroy_process_group()


def evaluate_once(params, wrapper, ids, fn, S, kw):
    return evalloop.evaluate(params, wrapper, ids, fn, S, **kw)


if __name__ == "__main__":
    main()
