"""Entry point mirroring the reference's `python main.py` (Experiments/Pythia-70M/main.py:23-32,
Experiments/Qwen2-0.5B/main.py:100-207): reads params.json, runs the sliding-window sweep, writes
the same pickle files.  One process per GPU (torchrun) shards the windows.

    python -m splitq.main --family pythia --params params.json [--synthetic-tokens 4096]

Offline there is no WikiText and no checkpoint: `--synthetic-tokens N` evaluates a seeded random
token stream on random-init weights of the named architecture (BASELINE.json: all runs synthetic).
"""
from __future__ import annotations

import argparse
import json
import os

import torch

from . import compat
from .evalloop import evaluate, make_importance_fn
from .models import Pythia70Model, QwenPointFiveBModel

DEFAULT_PARAMS = {
    "pythia": {"ratios": [0, 0.25, 0.5, 0.75, 1], "layers_of_interest": [1, 2, 4, 5], "stride": 32,
               "experiment": "last_row", "methods": ["regular_importance", "last_row", "aggregate_till"]},
    "qwen": {"layers_of_interest": [3, 11, 18, 22, 23], "methods": ["regular_importance", "last_row", "aggregate_till"],
             "max_length": 512, "stride": 32, "ratios": [0, 0.25, 0.5, 0.75, 1]},
}


def run(family: str, params: dict, n_tokens: int, device: str, seed: int = 0, tiny: bool = False,
        out_prefix=None, split_devices=None, log=print):
    """Build the wrapper, a synthetic token stream and run evaluate(); returns EvalResult."""
    if family == "pythia":
        if params.get("experiment", "last_row") not in ("last_row",):
            # Pythia-70M/main.py:28-32: "initial" is the reference's broken path
            raise ValueError("Invalid experiment type")
        cfg = compat.PYTHIA_TINY if tiny else compat.PYTHIA_70M
        wrapper = Pythia70Model.from_config(device, params["ratios"], cfg=cfg, seed=seed,
                                            scale_mode=params.get("scale_mode", "group"),
                                            split_devices=split_devices)
        # last_row_exp.py:72-73: window = model max positions, stride hard-coded to 32
        max_length = params.get("max_length", cfg["max_position_embeddings"])
        params = dict(params, stride=32)
    elif family == "qwen":
        cfg = compat.QWEN2_TINY if tiny else compat.QWEN2_05B
        wrapper = QwenPointFiveBModel.from_config(device, cfg=cfg, seed=seed,
                                                  scale_mode=params.get("scale_mode", "group"),
                                                  split_devices=split_devices)
        max_length = params["max_length"]
    else:
        raise ValueError(f"unknown family: {family}")
    gen = torch.Generator().manual_seed(1234)
    ids = torch.randint(0, cfg["vocab_size"], (1, n_tokens), generator=gen)
    head_w = params.get("head_weights")
    fn = make_importance_fn(wrapper, params["methods"], params.get("importance_impl", "fused"), head_w,
                            layers=params["layers_of_interest"])
    return evaluate(params, wrapper, ids, fn, max_length, out_prefix=out_prefix,
                    reduce_device=wrapper.device if torch.distributed.is_initialized() else None, log=log,
                    batch_methods=bool(params.get("batch_methods", False)),
                    batch_combos=bool(params.get("batch_combos", False)))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--family", choices=["pythia", "qwen"], required=True)
    ap.add_argument("--params", default=None, help="params.json with the reference's keys")
    ap.add_argument("--synthetic-tokens", type=int, default=4096)
    ap.add_argument("--out-prefix", default="")
    args = ap.parse_args()
    params = json.load(open(args.params)) if args.params else DEFAULT_PARAMS[args.family]
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    res = run(args.family, params, args.synthetic_tokens, f"cuda:{local}", out_prefix=args.out_prefix)
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"n_tokens": res.n_tokens, "windows": res.windows, "ppl": res.ppl}))
    if torch.distributed.is_initialized():
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
