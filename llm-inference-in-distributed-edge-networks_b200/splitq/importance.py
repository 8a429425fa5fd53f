"""`get_importance_order` with the reference's signature and return contract, computed by K1.

Reference: Experiments/Pythia-70M/last_row_exp.py:9-45 (three methods) and
Experiments/Qwen2-0.5B/main.py:21-98 (adds "weighted_importance" and the head_weights
argument).  Same method names, same list-of-(S,)-tensors result, same ValueError text.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from .boundary import importance_from_attention

METHODS = ("regular_importance", "weighted_importance", "last_row", "aggregate_till")


def get_importance_order(method: str, attention_map, num_layers: int, head_weights=None,
                         layers: Optional[Sequence[int]] = None) -> List[Optional[torch.Tensor]]:
    """attention_map[l]: (1,H,S,S) fp32 CUDA tensor.  Returns one (S,) tensor per layer.

    `layers` (extension): only score these layers (others are None).  The reference scores every
    layer although only `layers_of_interest` are used (last_row_exp.py:12,109); "aggregate_till"
    needs all layers up to the largest requested one regardless.
    """
    if method not in METHODS:
        raise ValueError(f"Unknown method: {method}")
    want = set(range(num_layers)) if layers is None else set(layers)
    res: List[Optional[torch.Tensor]] = []
    running = None
    for l in range(num_layers):
        need = l in want or (method == "aggregate_till" and l <= max(want, default=-1))
        if not need:
            res.append(None)
            continue
        A = attention_map[l]
        if method == "last_row":
            _, imp = importance_from_attention(A, want_col=False, want_last=True)
        elif method == "weighted_importance":
            # Qwen main.py:57-78: sum_h w[l][h] * A_h, then column mean (no division by H)
            imp, _ = importance_from_attention(A, head_w=[float(w) for w in head_weights[l]])
        else:
            imp, _ = importance_from_attention(A)
        imp = imp[0] if imp.shape[0] == 1 else imp
        if method == "aggregate_till":
            running = imp.clone() if running is None else running + imp  # last_row_exp.py:38-39
            imp = running / (l + 1)
        res.append(imp if l in want else None)
    return res
