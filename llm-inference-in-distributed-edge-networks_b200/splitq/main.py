"""Entry point mirroring the reference's `python main.py` (Experiments/Pythia-70M/main.py:23-32,
Experiments/Qwen2-0.5B/main.py:100-207): reads params.json, runs the sliding-window sweep, writes
the same pickle files.  One process per GPU (torchrun) shards the windows.

    python -m splitq.main --family pythia --params params.json [--synthetic-tokens 4096]
    torchrun --nproc-per-node 8 -m splitq.main --family qwen --pairs      # 4 GPU pairs

Offline there is no WikiText and no checkpoint: `--synthetic-tokens N` evaluates a seeded random
token stream on random-init weights of the named architecture (BASELINE.json: all runs synthetic).

params.json keys.  The reference's own: ratios, layers_of_interest, methods,
stride, experiment (Pythia), max_length (Qwen).  Optional, with defaults that reproduce the
reference: `scale_mode` ("group" | "token"), `bits` (width of the single rung, default 4) or
`ladder` ([[fraction, bits], ...] scaled by ratio, e.g. [[0.333, 2], [0.333, 4], [0.334, 8]]),
`importance_impl` ("fused" | "materialised"), `pairs` (true: GPU-pair layout, upstream half on rank
2p, downstream half on rank 2p+1), `split_devices` ([up, down] device names driven by ONE process),
`batch_windows` is accepted for forward compatibility (the sweep is batched over (method, ratio)
variants instead: `batch_methods`, `batch_combos`), `head_weights` (24 x 14 floats) or
`head_weights_file` (a pickle like the reference's ../../attention_head_weights.pkl, Qwen main.py:129).
"""
from __future__ import annotations

import argparse
import json
import os
import pickle

import torch

from . import compat
from .evalloop import FILE_NAMES, evaluate, evaluate_channel, make_importance_fn
from .models import Pythia70Model, QwenPointFiveBModel

DEFAULT_PARAMS = {
    "pythia": {"ratios": [0, 0.25, 0.5, 0.75, 1], "layers_of_interest": [1, 2, 4, 5], "stride": 32,
               "experiment": "last_row", "methods": ["regular_importance", "last_row", "aggregate_till"]},
    "qwen": {"layers_of_interest": [3, 11, 18, 22, 23], "methods": ["regular_importance", "last_row", "aggregate_till"],
             "max_length": 512, "stride": 32, "ratios": [0, 0.25, 0.5, 0.75, 1]},
}


def load_head_weights(params: dict, num_layers: int, num_heads: int):
    """`weighted_importance` needs per-layer, per-head weights: the reference unpickles
    ../../attention_head_weights.pkl (Qwen2-0.5B/main.py:129-130, the Relevance experiment's output)."""
    hw = params.get("head_weights")
    path = params.get("head_weights_file")
    if hw is None and path is None and os.path.exists("../../attention_head_weights.pkl"):
        path = "../../attention_head_weights.pkl"
    if hw is None and path is not None:
        with open(path, "rb") as f:
            hw = pickle.load(f)
    if hw is None:
        if "weighted_importance" in params.get("methods", []):
            raise FileNotFoundError("weighted_importance needs params['head_weights'] / ['head_weights_file'] "
                                    "(the reference reads ../../attention_head_weights.pkl)")
        return None
    hw = [[float(x) for x in row] for row in hw]
    if len(hw) != num_layers or any(len(row) != num_heads for row in hw):
        raise ValueError(f"head weights must be {num_layers} x {num_heads}")
    return hw


def run(family: str, params: dict, n_tokens: int, device: str, seed: int = 0, tiny: bool = False,
        out_prefix=None, split_devices=None, pairs=None, log=print):
    """Build the wrapper, a synthetic token stream and run evaluate(); returns EvalResult."""
    split_devices = split_devices if split_devices is not None else params.get("split_devices")
    pairs = bool(params.get("pairs", False)) if pairs is None else pairs
    if pairs and split_devices:
        raise ValueError("`pairs` (two processes per pair) and `split_devices` (one process, two devices) exclude each other")
    kw = dict(scale_mode=params.get("scale_mode", "group"), bits=int(params.get("bits", 4)),
              ladder_fracs=[(float(f), int(b)) for f, b in params["ladder"]] if params.get("ladder") else None,
              split_devices=split_devices)
    methods = params["methods"]
    channel = bool(methods) and "channel" in methods[0]  # Qwen2-0.5B/main.py:118-119 (intent of `.contains`)
    if family == "pythia":
        if params.get("experiment", "last_row") not in ("last_row",):
            # Pythia-70M/main.py:28-32: "initial" is the reference's broken path
            raise ValueError("Invalid experiment type")
        cfg = compat.PYTHIA_TINY if tiny else compat.PYTHIA_70M
        wrapper = Pythia70Model.from_config(device, params["ratios"], cfg=cfg, seed=seed, **kw)
        # last_row_exp.py:72-73: window = model max positions, stride hard-coded to 32
        max_length = params.get("max_length", cfg["max_position_embeddings"])
        params = dict(params, stride=32)
    elif family == "qwen":
        cfg = compat.QWEN2_TINY if tiny else compat.QWEN2_05B
        wrapper = QwenPointFiveBModel.from_config(device, cfg=cfg, seed=seed, **kw)
        max_length = params["max_length"]
    else:
        raise ValueError(f"unknown family: {family}")
    gen = torch.Generator().manual_seed(1234)
    ids = torch.randint(0, cfg["vocab_size"], (1, n_tokens), generator=gen)
    reduce_device = wrapper.device if torch.distributed.is_initialized() and \
        torch.distributed.get_backend() == "nccl" else None
    if channel:
        if family != "qwen":
            raise ValueError("channel-wise methods exist for the Qwen wrapper only (qwen_layer_wise.py:106)")
        # channel_wise.py:25-26 hard-codes the window and the stride
        return evaluate_channel(params["layers_of_interest"], methods, wrapper, ids, 512, 32, out_prefix=out_prefix,
                                reduce_device=reduce_device, log=log)
    head_w = load_head_weights(params, cfg["num_hidden_layers"], cfg["num_attention_heads"])
    fn = make_importance_fn(wrapper, methods, params.get("importance_impl", "fused"), head_w,
                            layers=params["layers_of_interest"])
    return evaluate(params, wrapper, ids, fn, max_length, out_prefix=out_prefix, reduce_device=reduce_device, log=log,
                    batch_methods=bool(params.get("batch_methods", False)),
                    batch_combos=bool(params.get("batch_combos", False)), pairs=pairs,
                    file_names=FILE_NAMES[family])


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--family", choices=["pythia", "qwen"], required=True)
    ap.add_argument("--params", default=None, help="params.json with the reference's keys")
    ap.add_argument("--synthetic-tokens", type=int, default=4096)
    ap.add_argument("--out-prefix", default="")
    ap.add_argument("--pairs", action="store_true", help="GPU-pair layout (same as \"pairs\": true in params.json)")
    args = ap.parse_args()
    params = json.load(open(args.params)) if args.params else DEFAULT_PARAMS[args.family]
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    res = run(args.family, params, args.synthetic_tokens, f"cuda:{local}", out_prefix=args.out_prefix,
              pairs=True if args.pairs else None)
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"n_tokens": res.n_tokens, "windows": res.windows, "ppl": res.ppl}))
    if torch.distributed.is_initialized():
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
