#!/usr/bin/env python
"""BASELINE.json configs[3]: layer-wise importance sweep over every split point of Qwen2-0.5B
("early vs final-layer importance correlation"), windows sharded over ranks.

For each synthetic window the fused kernel (through the HF attention hook) scores every layer in
ONE SDPA forward; per layer pair the Jensen-Shannon divergence of the column-mean importance
distributions (Notebooks/distributions_distance_across_layers.ipynb cells 12-16) and, per layer,
the Spearman correlation with the final layer's ranking are averaged over windows and all-reduced.

    python benchmarks/layer_sweep.py [--windows 16] [--S 4096]
    torchrun --nproc-per-node 8 benchmarks/layer_sweep.py --windows 64
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

from splitq import analytics, compat, hf_hook  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--family", default="qwen", choices=["qwen", "pythia"])
    ap.add_argument("--windows", type=int, default=16)
    ap.add_argument("--S", type=int, default=4096)
    args = ap.parse_args()
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=dev)
    family, cfg = ("qwen2", compat.QWEN2_05B) if args.family == "qwen" else ("gpt_neox", compat.PYTHIA_70M)
    model = compat.build_causal_lm(family, cfg, seed=0).to(dev)
    L = cfg["num_hidden_layers"]
    S = min(args.S, cfg["max_position_embeddings"])
    gen = torch.Generator().manual_seed(1234)
    ids_all = torch.randint(0, cfg["vocab_size"], (args.windows, S), generator=gen)
    mine = list(range(rank, args.windows, world))
    js_sum = torch.zeros(L, L, dtype=torch.float64, device=dev)
    sp_sum = torch.zeros(L, dtype=torch.float64, device=dev)
    hf_hook.fused_importance(model, ids_all[:1].to(dev), ["regular_importance"], L)  # warm-up
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for w in mine:
        imp = hf_hook.fused_importance(model, ids_all[w:w + 1].to(dev), ["regular_importance"], L)["regular_importance"]
        dist = torch.stack(imp).unsqueeze(1)                     # (L, 1, S); each row sums to 1
        js = analytics.layer_distance_matrix(dist)
        js_sum += torch.nan_to_num(js)
        sp_sum += analytics.spearman(dist[:, 0], dist[-1:, 0].expand(L, -1))
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    stats = torch.cat([js_sum.flatten(), sp_sum, torch.tensor([float(len(mine)), dt], dtype=torch.float64, device=dev)])
    if world > 1:
        tmax = torch.tensor([dt], dtype=torch.float64, device=dev)
        torch.distributed.all_reduce(stats[:-1], op=torch.distributed.ReduceOp.SUM)
        torch.distributed.all_reduce(tmax, op=torch.distributed.ReduceOp.MAX)
        dt = tmax.item()
    n = stats[-2].item()
    js = (stats[:L * L].view(L, L) / n).cpu()
    sp = (stats[L * L:L * L + L] / n).cpu()
    if rank == 0:
        print(json.dumps({
            "family": args.family, "n_gpus": world, "windows": int(n), "S": S, "layers": L, "seconds": dt,
            "windows_per_s": n / dt, "layer_scorings_per_s": n * L / dt, "tokens_scored_per_s": n * L * S / dt,
            "js_divergence_first_row": [round(x, 5) for x in js[0, 1:].tolist()],
            "js_divergence_adjacent": [round(js[i, i + 1].item(), 5) for i in range(L - 1)],
            "spearman_vs_final_layer": [round(x, 4) for x in sp.tolist()]}))
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
