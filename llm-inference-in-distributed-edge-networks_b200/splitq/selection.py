"""Selection / aggregation variants of the reference's first experiment.

Reference: `extract_attentions` (Experiments/Pythia-70M/initial_exp.py:27-72) builds token
orderings from the column-mean importance of a layer, of layers 0-2 averaged ("aggregate upto 2"),
of their element-wise maximum ("maximum aggregation"), or returns the raw layer-2 distribution
("upto ratio"); `batched_top_rho_quantization` (pythia_model.py:93-109) keeps the fewest tokens
whose importance mass reaches 1 - 0.1*ratio and quantises the rest.  The importance vectors come
from K1 (or K2); the bookkeeping on (S,) vectors below is plain torch on the GPU.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import numpy as np
import torch

from .boundary import importance_from_attention


def _col_importance(A: torch.Tensor) -> torch.Tensor:
    col, _ = importance_from_attention(A)
    return col[0]


def extract_attentions(attention_weights, interesting_layers) -> Dict[object, np.ndarray]:
    """Same keys and value types as initial_exp.py:27-72: ascending argsort orderings (numpy int64),
    except "upto ratio", which maps to the layer-2 importance distribution itself (numpy fp32)."""
    out: Dict[object, np.ndarray] = {}
    for layer in interesting_layers:
        if isinstance(layer, str) and layer == "aggregate upto 2":
            agg = sum(_col_importance(attention_weights[i]) for i in range(3)) / 3
            out[layer] = torch.argsort(agg, descending=False, stable=True).cpu().numpy()
        elif layer == "maximum aggregation":
            agg = torch.zeros_like(_col_importance(attention_weights[0]))
            for i in range(3):
                agg = torch.maximum(agg, _col_importance(attention_weights[i]))
            out[layer] = torch.argsort(agg, descending=False, stable=True).cpu().numpy()
        elif layer == "upto ratio":
            out[layer] = _col_importance(attention_weights[2]).cpu().numpy()
        else:
            out[layer] = torch.argsort(_col_importance(attention_weights[layer]), descending=False, stable=True).cpu().numpy()
    return out


def top_rho_keep_count(distribution: torch.Tensor, ratio: float) -> int:
    """Fewest most-important tokens whose importance sums to >= 1 - 0.1*ratio
    (pythia_model.py:93-106: sort descending, accumulate until the threshold is met)."""
    threshold = 1 - 0.1 * ratio
    vals = torch.sort(distribution.double(), descending=True).values
    before = torch.cumsum(vals, dim=0) - vals  # running total before each token is taken
    return int((before < threshold).sum().item())  # the loop takes a token while total < threshold


def top_rho_ladder(distribution: torch.Tensor, ratio: float, bits: int = 4) -> List[Tuple[int, int]]:
    """Ladder for select_bits: everything outside the kept top-rho set is quantised (the intended
    semantics of batched_top_rho_quantization; its own slice mixes token ids and ranks)."""
    S = distribution.numel()
    k = S - top_rho_keep_count(distribution, ratio)
    return [(k, bits)] if k > 0 else []


def quantization_sets_from_ordering(ordering: Sequence[int], ratios: Sequence[int]) -> List[np.ndarray]:
    """tokens_to_quantize per ratio as in batched_quantization_helper (pythia_model.py:141-142):
    the first int(0.1 * ratio * len(ordering)) entries of the ascending ordering."""
    ordering = np.asarray(ordering)
    return [ordering[: int(0.1 * r * len(ordering))] for r in ratios]
