"""Sliding-window perplexity evaluation over (method x split layer x ratio), data-parallel.

Reference loop: Experiments/Pythia-70M/last_row_exp.py:85-143 and Experiments/Qwen2-0.5B/main.py:
151-207 -- windows [b, min(b+max_length, N)), labels -100 on the context part, num_loss_tokens =
valid - batch, total_nll[m][l][r] += nll * num_loss_tokens, ppl = exp(total / n_tokens), pickle
dumps every 1000 windows and at the end (same file names: FILE_NAMES).

What changes here, none of it numerically:
  * windows are independent (prev_end_loc is a closed form of the window index), so rank r of W
    takes windows k = r (mod W) -- or, with pairs=True, GPU pair p of P takes k = p (mod P), the
    upstream half on rank 2p and the downstream half on rank 2p+1 with the payload crossing between
    them; NLL sums (fp64) and token counts are all-reduced;
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
    hop_bytes: int = 0  # bytes this rank sent across the pair boundary (pairs=True, upstream ranks)


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


class PairLink:
    """The device hop between the two ranks of a GPU pair: rank 2p runs the upstream half
    on its GPU, rank 2p+1 the downstream half on its own; packed payloads (and, for ratio 0, the raw
    hidden states -- the reference skips quantisation there, pythia_model.py:167) cross with
    torch.distributed send/recv: NCCL over NVLink for CUDA tensors, staged through host memory for
    the gloo backend (CPU tests)."""

    def __init__(self, d: "_Dist") -> None:
        self.dist, self.group, self.peer = d.dist, d.group, d.rank ^ 1
        self.cpu_only = d.dist.get_backend(d.group) == "gloo"
        self.sent_bytes = 0

    def send(self, t: torch.Tensor) -> None:
        self.sent_bytes += t.numel() * t.element_size()
        self.dist.send(t.cpu().contiguous() if self.cpu_only else t.contiguous(), self.peer, group=self.group)

    def recv(self, like: torch.Tensor) -> torch.Tensor:
        if self.cpu_only and like.is_cuda:
            host = torch.empty(like.shape, dtype=like.dtype)
            self.dist.recv(host, self.peer, group=self.group)
            like.copy_(host)
        else:
            self.dist.recv(like, self.peer, group=self.group)
        return like


FILE_NAMES = {
    # Experiments/Qwen2-0.5B/main.py:189-191,206
    "qwen": ("total_nll.pkl", "n_tokens.pkl", "avg_ppl_results.pkl"),
    # Experiments/Pythia-70M/last_row_exp.py:122-124,142
    "pythia": ("total_nll_pythia_70m.pkl", "n_tokens_pythia_70m.pkl", "avg_ppl_results_pythia_70m.pkl"),
    # Experiments/Qwen2-0.5B/channel_wise.py:60-62,77
    "channel": ("total_nll_channel.pkl", "n_tokens_channel.pkl", "ppl_channel.pkl"),
}


def evaluate(params: dict, wrapper, input_ids: torch.Tensor, importance_fn: Callable, max_length: int,
             out_prefix: Optional[str] = None, checkpoint_every: int = 1000, reduce_device=None, group=None,
             log: Callable = print, batch_methods: bool = False, batch_combos: bool = False,
             pairs: bool = False, file_names: Sequence[str] = FILE_NAMES["qwen"]) -> EvalResult:
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
    pairs          GPU-pair layout of SURVEY 8e / BASELINE.json north_star: ranks (2p, 2p+1) form pair
                   p; rank 2p holds the upstream half (scoring, layers 0..i, select, quantise, pack),
                   rank 2p+1 the downstream half (dequantise, layers i+1.., loss); the payload crosses
                   between the two processes (PairLink); pair p of P = world // 2 owns windows
                   k = p (mod P).  With one rank (or pairs=False) both halves run in this process and
                   rank r of W owns windows k = r (mod W).
    out_prefix     if set, rank 0 writes <prefix> + file_names = (total_nll, n_tokens, avg_ppl) pickles
    """
    ratios, lois, methods = params["ratios"], params["layers_of_interest"], params["methods"]
    stride = params.get("stride", 32)
    d = _Dist(group)
    sched = window_schedule(input_ids.size(1), max_length, stride)
    M, L, R = len(methods), len(lois), len(ratios)
    if pairs and d.world >= 2:
        shards, shard = d.world // 2, d.rank // 2
        role = "idle" if d.rank >= 2 * shards else ("up" if d.rank % 2 == 0 else "down")
        link = PairLink(d) if role != "idle" else None
    else:
        shards, shard, role, link = d.world, d.rank, "both", None
    up, down = role in ("both", "up"), role in ("both", "down")
    pos_r = [r for r, ratio in enumerate(ratios) if ratio > 0]
    zero_r = [r for r, ratio in enumerate(ratios) if not ratio > 0]

    def cross_payload(hid, impv, ratio, rows, S):
        """quantise + pack upstream, hop, -> the payload on the downstream side (None on an upstream rank)"""
        payload = wrapper.boundary_send(hid, impv, ratio) if up else None
        if role == "both":
            return wrapper.boundary_hop(payload)
        if role == "up":
            link.send(payload)
            return None
        return link.recv(wrapper.payload_like(rows, S, ratio))

    def cross_hidden(hidden, S):
        """ratio 0: the reference skips the quantiser (pythia_model.py:167), the fp32 states cross as they are"""
        if role == "both":
            return hidden
        if role == "up":
            link.send(hidden)
            return None
        return link.recv(wrapper.hidden_like(1, S))

    total = torch.zeros(M, L, R, dtype=torch.float64)
    n_tokens = 0
    done = 0
    full_rounds = len(sched) // shards  # rounds of `shards` consecutive windows in which every shard has one
    for k, (begin, end, trg_len) in enumerate(sched):
        if role == "idle" or k % shards != shard:
            continue
        ids = input_ids[:, begin:end].to(wrapper.device)
        target = ids.clone()
        target[:, :-trg_len] = -100
        num_loss_tokens = int((target != -100).sum().item()) - target.size(0)
        imp = importance_fn(ids) if up else None
        shape = tuple(ids.shape)
        S = shape[1]
        with torch.no_grad():
            nll_dev = torch.zeros(M, L, R, dtype=torch.float32, device=wrapper.down_device) if down else None
            for l, loi in enumerate(lois):
                hidden = wrapper.upstream(ids, loi) if up else None  # once per split layer
                combos = batch_combos and shape[0] == 1
                by_method = batch_methods and M > 1 and shape[0] == 1 and not combos
                hid_m = imp_m = None
                if up and (combos or by_method):
                    hid_m = hidden.expand(M, -1, -1).contiguous()
                    imp_m = torch.stack([imp[method][loi] for method in methods])
                if combos:
                    payloads = [cross_payload(hid_m, imp_m, ratios[r], M, S) for r in pos_r]
                    if down and pos_r:
                        rows = M * len(pos_r)
                        nll = wrapper.downstream((rows, S), target.expand(rows, -1), loi, payload=payloads,
                                                 per_sequence=True).view(len(pos_r), M)
                        for j, r in enumerate(pos_r):
                            nll_dev[:, l, r] = nll[j]
                elif by_method:
                    # one downstream pass per ratio with the M methods as the batch
                    for r in pos_r:
                        payload = cross_payload(hid_m, imp_m, ratios[r], M, S)
                        if down:
                            nll_dev[:, l, r] = wrapper.downstream((M, S), target.expand(M, -1), loi, payload=payload,
                                                                  per_sequence=True)
                else:
                    for m, method in enumerate(methods):
                        for r in pos_r:
                            payload = cross_payload(hidden, imp[method][loi] if up else None, ratios[r], shape[0], S)
                            if down:
                                nll_dev[m, l, r] = wrapper.downstream(shape, target, loi, payload=payload)
                if zero_r:  # ratio 0 does not depend on the method (reference: same NLL in every row)
                    raw = cross_hidden(hidden, S)
                    if down:
                        base = wrapper.downstream(shape, target, loi, hidden=raw)
                        for r in zero_r:
                            nll_dev[:, l, r] = base
        if down:
            total += nll_dev.double().cpu() * num_loss_tokens  # one sync per window instead of M*L*R
            n_tokens += num_loss_tokens
            done += 1
        # every rank reaches the collective: the gate depends only on the global window index
        rnd = k // shards
        if out_prefix is not None and checkpoint_every and (rnd + 1) % checkpoint_every == 0 and rnd < full_rounds:
            _checkpoint(d, total.clone(), n_tokens, out_prefix, reduce_device, log, rnd + 1, file_names)
    if role == "idle" and out_prefix is not None and checkpoint_every:
        for rnd in range(full_rounds):  # an unpaired last rank still takes part in the collectives
            if (rnd + 1) % checkpoint_every == 0:
                _checkpoint(d, total.clone(), 0, out_prefix, reduce_device, log, rnd + 1, file_names)
    # final reduction (upstream and idle ranks contribute zeros: the sums live on the downstream ranks)
    red = torch.cat([total.flatten(), torch.tensor([float(n_tokens), float(done)], dtype=torch.float64)])
    if reduce_device is not None:
        red = red.to(reduce_device)
    red = d.allreduce_sum(red).cpu()
    total = red[:-2].view(M, L, R)
    n_all, w_all = int(round(red[-2].item())), int(round(red[-1].item()))
    res = EvalResult(total.tolist(), n_all, w_all)
    res.ppl = [[[math.exp(total[m, l, r].item() / n_all) if n_all else float("nan") for r in range(R)]
                for l in range(L)] for m in range(M)]
    res.hop_bytes = link.sent_bytes if link is not None else 0
    if out_prefix is not None and d.rank == 0:
        with open(out_prefix + file_names[0], "wb") as f:
            pickle.dump(res.total_nll, f)
        with open(out_prefix + file_names[1], "wb") as f:
            pickle.dump(res.n_tokens, f)
        with open(out_prefix + file_names[2], "wb") as f:
            pickle.dump(res.ppl, f)
        log("Average ppl results", res.ppl)
    return res


def _checkpoint(d: _Dist, total, n_tokens, out_prefix, reduce_device, log, done, file_names=FILE_NAMES["qwen"]):
    red = torch.cat([total.flatten(), torch.tensor([float(n_tokens)], dtype=torch.float64)])
    if reduce_device is not None:
        red = red.to(reduce_device)
    red = d.allreduce_sum(red).cpu()
    if d.rank == 0:
        log(f"Processed {done} chunks per shard")
        with open(out_prefix + file_names[0], "wb") as f:
            pickle.dump(red[:-1].view(total.shape).tolist(), f)
        with open(out_prefix + file_names[1], "wb") as f:
            pickle.dump(int(round(red[-1].item())), f)


def evaluate_channel(layers_of_interest: Sequence[int], methods: Sequence[str], wrapper, input_ids: torch.Tensor,
                     max_length: int = 512, stride: int = 32, out_prefix: Optional[str] = None, reduce_device=None,
                     group=None, log: Callable = print) -> EvalResult:
    """The channel-wise sweep of Experiments/Qwen2-0.5B/channel_wise.py:10-78 (max_length 512 and
    stride 32 are hard-coded there, :25-26): total_nll[l][m] += NLL(channel_wise(ids, target, layer,
    method)) * num_loss_tokens over the sliding windows, ppl = exp(total / n_tokens), written to
    total_nll_channel.pkl / n_tokens_channel.pkl / ppl_channel.pkl.  Windows are sharded over ranks
    and the upstream half runs once per (window, layer) instead of once per method; the sums are
    all-reduced in fp64.  EvalResult.total_nll / .ppl are indexed [layer][method] like the reference's
    lists."""
    from . import boundary

    d = _Dist(group)
    sched = window_schedule(input_ids.size(1), max_length, stride)
    L, M = len(layers_of_interest), len(methods)
    for method in methods:
        if method not in ("channel_8", "channel_4", "channel_1_mean", "channel_1_max"):
            raise ValueError(f"Unknown method: {method}")  # qwen_layer_wise.py:152
    total = torch.zeros(L, M, dtype=torch.float64)
    n_tokens = done = 0
    for k, (begin, end, trg_len) in enumerate(sched):
        if k % d.world != d.rank:
            continue
        ids = input_ids[:, begin:end].to(wrapper.device)
        target = ids.clone()
        target[:, :-trg_len] = -100
        num_loss_tokens = int((target != -100).sum().item()) - target.size(0)
        with torch.no_grad():
            nll_dev = torch.zeros(L, M, dtype=torch.float32, device=wrapper.down_device)
            for l, loi in enumerate(layers_of_interest):
                hidden = wrapper.upstream(ids, loi)
                for m, method in enumerate(methods):
                    quantised, _, _ = boundary.channel_fake_quant(hidden, method)
                    nll_dev[l, m] = wrapper.downstream(tuple(ids.shape), target, loi, hidden=quantised)
        total += nll_dev.double().cpu() * num_loss_tokens
        n_tokens += num_loss_tokens
        done += 1
    red = torch.cat([total.flatten(), torch.tensor([float(n_tokens), float(done)], dtype=torch.float64)])
    if reduce_device is not None:
        red = red.to(reduce_device)
    red = d.allreduce_sum(red).cpu()
    total = red[:-2].view(L, M)
    n_all, w_all = int(round(red[-2].item())), int(round(red[-1].item()))
    res = EvalResult(total.tolist(), n_all, w_all)
    res.ppl = [[math.exp(total[l, m].item() / n_all) if n_all else float("nan") for m in range(M)] for l in range(L)]
    if out_prefix is not None and d.rank == 0:
        names = FILE_NAMES["channel"]
        for name, obj in zip(names, (res.total_nll, res.n_tokens, res.ppl)):
            with open(out_prefix + name, "wb") as f:
                pickle.dump(obj, f)
        log(res.ppl)
    return res


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
