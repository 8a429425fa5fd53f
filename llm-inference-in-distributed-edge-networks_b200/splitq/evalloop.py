"""Sliding-window perplexity evaluation over (method x split layer x ratio), data-parallel.

Reference loop: Experiments/Pythia-70M/last_row_exp.py:85-143 and Experiments/Qwen2-0.5B/main.py:
151-207 -- windows [b, min(b+max_length, N)), labels -100 on the context part, num_loss_tokens =
valid - batch, total_nll[m][l][r] += nll * num_loss_tokens, ppl = exp(total / n_tokens), pickle
dumps every 1000 windows and at the end (same file names).

What changes here, none of it numerically:
  * windows are independent (prev_end_loc is a closed form of the window index), so rank r of W
    takes windows k = r (mod W); NLL sums (fp64) and token counts (int64) are all-reduced;
  * the upstream half runs once per split layer and window, and its hidden states are reused for
    every (method, ratio) -- the reference re-runs the whole model M*L*R times per window;
  * importance comes either from the fused Q/K kernel through the HF attention hook ("fused") or
    from materialised eager attention maps ("materialised", the reference's way).
"""
from __future__ import annotations

import math
import pickle
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence

import torch


def window_schedule(n_tokens: int, max_length: int, stride: int):
    """[(begin, end, trg_len)] of last_row_exp.py:85-90,115,129-130, in closed form."""
    out = []
    k = 0
    while True:
        begin = k * stride
        if begin >= n_tokens:
            break
        end = min(begin + max_length, n_tokens)
        prev_end = 0 if k == 0 else min(begin - stride + max_length, n_tokens)
        out.append((begin, end, end - prev_end))
        if end == n_tokens:
            break
        k += 1
    return out


@dataclass
class EvalResult:
    total_nll: List[List[List[float]]]
    n_tokens: int
    windows: int
    ppl: List[List[List[float]]] = field(default_factory=list)


class _Dist:
    """Thin view of torch.distributed that degrades to a single process."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist = dist
        self.on = dist.is_available() and dist.is_initialized()
        self.group = group
        self.rank = dist.get_rank(group) if self.on else 0
        self.world = dist.get_world_size(group) if self.on else 1

    def allreduce_sum(self, t: torch.Tensor) -> torch.Tensor:
        if self.on and self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        return t


def evaluate(params: dict, wrapper, input_ids: torch.Tensor, importance_fn: Callable, max_length: int,
             out_prefix: Optional[str] = None, checkpoint_every: int = 1000, reduce_device=None, group=None,
             log: Callable = print, batch_methods: bool = False, batch_combos: bool = False) -> EvalResult:
    """Run the (method, layer, ratio) sweep over this rank's share of the windows.

    params         the reference's params.json dict: ratios, layers_of_interest, methods, stride
                   (last_row_exp.py:59-64; Qwen main.py:112-116,137-138)
    wrapper        a splitq.models wrapper (needs upstream/boundary_send/boundary_hop/downstream)
    input_ids      (1, N) token stream (encodings.input_ids)
    importance_fn  (window_ids (1,S)) -> {method: [importance (S,) per layer]}
    batch_methods  push the M methods of one (layer, ratio) through the downstream half as one batch
                   (same arithmetic per row; cuBLAS may pick other kernels, so NLLs agree to ~1e-6)
    batch_combos   push ALL (method, ratio > 0) variants of one split layer through the downstream half
                   as one batch (the payloads of different ratios have different widths; they are
                   dequantised one by one and their rows concatenated): M*(R-1) rows per forward
    out_prefix     if set, rank 0 writes <prefix>total_nll*.pkl / n_tokens*.pkl / avg_ppl_results*.pkl
    """
    ratios, lois, methods = params["ratios"], params["layers_of_interest"], params["methods"]
    stride = params.get("stride", 32)
    d = _Dist(group)
    sched = window_schedule(input_ids.size(1), max_length, stride)
    M, L, R = len(methods), len(lois), len(ratios)
    total = torch.zeros(M, L, R, dtype=torch.float64)
    n_tokens = 0
    done = 0
    for k, (begin, end, trg_len) in enumerate(sched):
        if k % d.world != d.rank:
            continue
        ids = input_ids[:, begin:end].to(wrapper.device)
        target = ids.clone()
        target[:, :-trg_len] = -100
        num_loss_tokens = int((target != -100).sum().item()) - target.size(0)
        imp = importance_fn(ids)
        with torch.no_grad():
            nll_dev = torch.zeros(M, L, R, dtype=torch.float32, device=wrapper.down_device)
            for l, loi in enumerate(lois):
                hidden = wrapper.upstream(ids, loi)  # once per split layer
                shape = tuple(ids.shape)
                base = None
                if batch_combos and hidden.shape[0] == 1:
                    hid_m = hidden.expand(M, -1, -1).contiguous()
                    imp_m = torch.stack([imp[method][loi] for method in methods])
                    pos_r = [r for r, ratio in enumerate(ratios) if ratio > 0]
                    if pos_r:
                        payloads = [wrapper.boundary_hop(wrapper.boundary_send(hid_m, imp_m, ratios[r])) for r in pos_r]
                        rows = M * len(pos_r)
                        nll = wrapper.downstream((rows, shape[1]), target.expand(rows, -1), loi, payload=payloads,
                                                 per_sequence=True).view(len(pos_r), M)
                        for j, r in enumerate(pos_r):
                            nll_dev[:, l, r] = nll[j]
                    if len(pos_r) < R:  # ratio 0 does not depend on the method
                        base = wrapper.downstream(shape, target, loi, hidden=hidden)
                        for r, ratio in enumerate(ratios):
                            if not ratio > 0:
                                nll_dev[:, l, r] = base
                    continue
                if batch_methods and M > 1 and hidden.shape[0] == 1:
                    # one downstream pass per ratio with the M methods as the batch
                    hid_m = hidden.expand(M, -1, -1).contiguous()
                    tgt_m = target.expand(M, -1)
                    imp_m = torch.stack([imp[method][loi] for method in methods])
                    for r, ratio in enumerate(ratios):
                        if ratio > 0:
                            payload = wrapper.boundary_hop(wrapper.boundary_send(hid_m, imp_m, ratio))
                            nll_dev[:, l, r] = wrapper.downstream((M, shape[1]), tgt_m, loi, payload=payload,
                                                                  per_sequence=True)
                        else:
                            nll_dev[:, l, r] = wrapper.downstream(shape, target, loi, hidden=hidden)
                    continue
                for m, method in enumerate(methods):
                    for r, ratio in enumerate(ratios):
                        if ratio > 0:
                            payload = wrapper.boundary_hop(wrapper.boundary_send(hidden, imp[method][loi], ratio))
                            nll_dev[m, l, r] = wrapper.downstream(shape, target, loi, payload=payload)
                        else:
                            if base is None:  # ratio 0 does not depend on the method (reference: same NLL)
                                base = wrapper.downstream(shape, target, loi, hidden=hidden)
                            nll_dev[m, l, r] = base
        total += nll_dev.double().cpu() * num_loss_tokens  # one sync per window instead of M*L*R
        n_tokens += num_loss_tokens
        done += 1
        if out_prefix is not None and checkpoint_every and done % checkpoint_every == 0:
            _checkpoint(d, total.clone(), n_tokens, out_prefix, reduce_device, log, done)
    # final reduction
    red = torch.cat([total.flatten(), torch.tensor([float(n_tokens), float(done)], dtype=torch.float64)])
    if reduce_device is not None:
        red = red.to(reduce_device)
    red = d.allreduce_sum(red).cpu()
    total = red[:-2].view(M, L, R)
    n_all, w_all = int(round(red[-2].item())), int(round(red[-1].item()))
    res = EvalResult(total.tolist(), n_all, w_all)
    res.ppl = [[[math.exp(total[m, l, r].item() / n_all) if n_all else float("nan") for r in range(R)]
                for l in range(L)] for m in range(M)]
    if out_prefix is not None and d.rank == 0:
        with open(out_prefix + "total_nll.pkl", "wb") as f:
            pickle.dump(res.total_nll, f)
        with open(out_prefix + "n_tokens.pkl", "wb") as f:
            pickle.dump(res.n_tokens, f)
        with open(out_prefix + "avg_ppl_results.pkl", "wb") as f:
            pickle.dump(res.ppl, f)
        log("Average ppl results", res.ppl)
    return res


def _checkpoint(d: _Dist, total, n_tokens, out_prefix, reduce_device, log, done):
    red = torch.cat([total.flatten(), torch.tensor([float(n_tokens)], dtype=torch.float64)])
    if reduce_device is not None:
        red = red.to(reduce_device)
    red = d.allreduce_sum(red).cpu()
    if d.rank == 0:
        log(f"Processed {done} chunks per rank")
        with open(out_prefix + "total_nll.pkl", "wb") as f:
            pickle.dump(red[:-1].view(total.shape).tolist(), f)
        with open(out_prefix + "n_tokens.pkl", "wb") as f:
            pickle.dump(int(round(red[-1].item())), f)


def make_importance_fn(wrapper, methods: Sequence[str], impl: str = "fused", head_weights=None,
                       layers: Optional[Sequence[int]] = None, eager_model=None) -> Callable:
    """importance_fn for evaluate().  "fused": K2 through the HF attention hook on the wrapper's
    own SDPA model (no second copy).  "materialised": the reference's route -- an eager copy with
    output_attentions=True, reduced by K1."""
    num_layers = wrapper.num_layers
    if impl == "fused":
        from . import hf_hook

        def fn(ids):
            return hf_hook.fused_importance(wrapper.model, ids, methods, num_layers, head_weights, layers)
        return fn
    if impl == "materialised":
        from . import compat
        from .importance import get_importance_order
        eager = eager_model or compat.clone_with_attention(wrapper.model, "eager").to(wrapper.device)

        def fn(ids):
            with torch.no_grad():
                att = eager(input_ids=ids, output_attentions=True).attentions
            return {m: get_importance_order(m, att, num_layers, head_weights, layers) for m in methods}
        return fn
    raise ValueError(f"unknown importance_impl: {impl}")
