"""Selection / aggregation variants of the reference's first experiment.

Reference: `extract_attentions` (Experiments/Pythia-70M/initial_exp.py:27-72) builds token
orderings from the column-mean importance of a layer, of layers 0-2 averaged ("aggregate upto 2"),
of their element-wise maximum ("maximum aggregation"), or returns the raw layer-2 distribution
("upto ratio"); `batched_top_rho_quantization` (pythia_model.py:93-109) keeps the fewest tokens
whose importance mass reaches 1 - 0.1*ratio and quantises the rest.  The importance vectors come
from K1 (or K2); the top-rho threshold is a mode of the selection kernel
(splitq_select_bits_per_window), the orderings are torch argsorts of the kernel's scores.
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


def top_rho_thresholds(ratios: Sequence[float]) -> List[float]:
    """pythia_model.py:95: the mass the kept (unquantised) tokens must reach for each ratio."""
    return [1 - 0.1 * float(r) for r in ratios]


def top_rho_select(distribution: torch.Tensor, ratios: Sequence[float], D: int, bits: int = 4, want_counts: bool = False):
    """K3a's cumulative-mass mode for a batch: row b of `distribution` (B,S) -- or the one (S,)
    distribution shared by all rows -- keeps the fewest most-important tokens whose importance sums to
    >= 1 - 0.1*ratios[b] and quantises the rest to `bits` bits (the intended semantics of
    batched_top_rho_quantization, pythia_model.py:93-109; its own slice mixes token ids and ranks,
    SURVEY 8f-2).  Everything stays on the device: -> (bits, row_off[, quantised counts])."""
    from .boundary import select_bits_per_window
    if distribution.dim() == 1:
        distribution = distribution.unsqueeze(0).expand(len(ratios), -1)
    return select_bits_per_window(distribution.contiguous(), bits, D, mass_thresholds=top_rho_thresholds(ratios),
                                  want_counts=want_counts)


def top_rho_keep_count(distribution: torch.Tensor, ratio: float) -> int:
    """Fewest most-important tokens whose importance sums to >= 1 - 0.1*ratio (pythia_model.py:93-106),
    computed by the selection kernel; reads one integer back from the device."""
    _, _, k = top_rho_select(distribution.reshape(1, -1), [ratio], D=4, want_counts=True)
    return distribution.numel() - int(k.item())


def top_rho_ladder(distribution: torch.Tensor, ratio: float, bits: int = 4) -> List[Tuple[int, int]]:
    """Ladder for select_bits: everything outside the kept top-rho set is quantised."""
    S = distribution.numel()
    k = S - top_rho_keep_count(distribution, ratio)
    return [(k, bits)] if k > 0 else []


def quantization_sets_from_ordering(ordering: Sequence[int], ratios: Sequence[int]) -> List[np.ndarray]:
    """tokens_to_quantize per ratio as in batched_quantization_helper (pythia_model.py:141-142):
    the first int(0.1 * ratio * len(ordering)) entries of the ascending ordering."""
    ordering = np.asarray(ordering)
    return [ordering[: int(0.1 * r * len(ordering))] for r in ratios]
