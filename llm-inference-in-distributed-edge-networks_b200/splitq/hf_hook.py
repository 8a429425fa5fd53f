"""Importance scoring hooked into HF attention: no eager copy, no attention maps.

The reference keeps a second, eager copy of the model only to obtain `attentions`
(Experiments/Pythia-70M/last_row_exp.py:68-70,92-95; Experiments/Qwen2-0.5B/main.py:132-134,
158-160) and reduces them with get_importance_order.  Here an attention function registered with
transformers' AttentionInterface receives each layer's post-RoPE query/key states, hands the
attention output to the stock SDPA path unchanged, and -- when a capture is active -- runs the
fused K2 kernel on (query, key) to get the same per-token scores.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch

from . import boundary

NAME = "splitq_importance"
_registered = False


def register() -> str:
    """Idempotently register the attention function; returns its name."""
    global _registered
    if not _registered:
        from transformers import AttentionInterface
        from transformers.integrations.sdpa_attention import sdpa_attention_forward

        def splitq_attention(module, query, key, value, attention_mask, scaling=None, dropout=0.0, **kwargs):
            cap = getattr(module, "_splitq_capture", None)
            if cap is not None:
                cap.observe(module, query, key, scaling if scaling is not None else query.shape[-1] ** -0.5)
            return sdpa_attention_forward(module, query, key, value, attention_mask, dropout=dropout,
                                          scaling=scaling, **kwargs)

        AttentionInterface.register(NAME, splitq_attention)
        try:  # transformers >= 4.53: the causal-mask builder is looked up under the same name
            from transformers.masking_utils import AttentionMaskInterface, sdpa_mask
            AttentionMaskInterface.register(NAME, sdpa_mask)
        except ImportError:
            pass
        _registered = True
    return NAME


def _attention_modules(model):
    mods = []
    for layer in model.base_model.layers:
        mods.append(getattr(layer, "self_attn", None) or getattr(layer, "attention"))
    return mods


class ImportanceCapture:
    """Context manager: while active, every attention layer of `model` records column-mean and
    last-row importance of its (B,H,S,S) attention without materialising it."""

    def __init__(self, model, layers: Optional[Sequence[int]] = None, want_col: bool = True,
                 want_last: bool = True, head_weights=None, mode: int = boundary.N.FUSED_F16X3):
        self.model = model
        self.layers = None if layers is None else set(layers)
        self.want_col, self.want_last, self.head_weights, self.mode = want_col, want_last, head_weights, mode
        self.col: Dict[int, torch.Tensor] = {}
        self.last: Dict[int, torch.Tensor] = {}
        self.weighted: Dict[int, torch.Tensor] = {}

    def observe(self, module, query, key, scaling):
        l = module.layer_idx
        if self.layers is not None and l not in self.layers:
            return
        q, k = query.float(), key.float()
        # one scoring pass: the weighted column mean is a second weighting of the same per-head sums
        hw2 = [float(x) for x in self.head_weights[l]] if self.head_weights is not None else None
        res = boundary.importance_fused(q, k, scaling, causal=True, mode=self.mode, want_col=self.want_col or hw2 is not None,
                                        want_last=self.want_last, head_w2=hw2)
        col, last = res[0], res[1]
        if col is not None and self.want_col:
            self.col[l] = col
        if last is not None:
            self.last[l] = last
        if hw2 is not None:
            self.weighted[l] = res[2]

    def __enter__(self):
        register()
        self._prev = self.model.config._attn_implementation
        self.model.config._attn_implementation = NAME
        for m in _attention_modules(self.model):
            m._splitq_capture = self
        return self

    def __exit__(self, *exc):
        self.model.config._attn_implementation = self._prev
        for m in _attention_modules(self.model):
            m._splitq_capture = None
        return False


def fused_importance(model, input_ids, methods: Sequence[str], num_layers: int, head_weights=None,
                     layers: Optional[Sequence[int]] = None) -> Dict[str, List[Optional[torch.Tensor]]]:
    """One SDPA forward of `model` -> {method: [ (S,) importance per layer ]} with the same method
    names and semantics as get_importance_order (last_row_exp.py:9-45, Qwen main.py:21-98)."""
    for m in methods:
        if m not in ("regular_importance", "weighted_importance", "last_row", "aggregate_till"):
            raise ValueError(f"Unknown method: {m}")
    need_all = "aggregate_till" in methods
    cap_layers = None if (layers is None or need_all) else layers
    hw = head_weights if "weighted_importance" in methods else None
    with torch.no_grad(), ImportanceCapture(model, cap_layers, want_col=True, want_last="last_row" in methods,
                                            head_weights=hw) as cap:
        model.base_model(input_ids=input_ids)
    want = set(range(num_layers)) if layers is None else set(layers)
    out: Dict[str, List[Optional[torch.Tensor]]] = {}
    squeeze = lambda t: t[0] if t.shape[0] == 1 else t  # noqa: E731
    for m in methods:
        res: List[Optional[torch.Tensor]] = []
        running = None
        for l in range(num_layers):
            if m == "aggregate_till":
                if l in cap.col:
                    running = cap.col[l].clone() if running is None else running + cap.col[l]
                res.append(squeeze(running / (l + 1)) if (l in want and running is not None) else None)
                continue
            src = {"regular_importance": cap.col, "last_row": cap.last, "weighted_importance": cap.weighted}[m]
            res.append(squeeze(src[l]) if (l in want and l in src) else None)
        out[m] = res
    return out
