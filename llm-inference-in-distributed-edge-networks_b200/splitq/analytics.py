"""Layer-to-layer comparison of importance distributions.

Reference: Notebooks/distributions_distance_across_layers.ipynb cells 12, 13 and 16 -- KL divergence
in bits with a p != 0 guard, Jensen-Shannon divergence, and the upper-triangular matrix of JS
divergences between per-layer column-mean importance distributions averaged over samples (the
"early vs final layer importance correlation" read-out of BASELINE.json configs[3]).  Spearman rank
correlation is added because the question being asked is about orderings.  Pure torch on whatever
device the (L, N, S) importance tensor lives on -- the fused kernel already produces it on the GPU.
"""
from __future__ import annotations

import torch


def kl_divergence(p: torch.Tensor, q: torch.Tensor) -> torch.Tensor:
    """sum(where(p != 0, p * log2(p / q), 0)) over the last dim (notebook cell 12)."""
    safe_p = torch.where(p != 0, p, torch.ones_like(p))
    return torch.where(p != 0, p * torch.log2(safe_p / q), torch.zeros_like(p)).sum(dim=-1)


def jensen_shannon(p: torch.Tensor, q: torch.Tensor) -> torch.Tensor:
    """0.5 * (KL(p||m) + KL(q||m)), m = (p+q)/2 (notebook cell 13, `jenson_shannon_distance`)."""
    m = 0.5 * (p + q)
    return 0.5 * (kl_divergence(p, m) + kl_divergence(q, m))


def layer_distance_matrix(dist: torch.Tensor) -> torch.Tensor:
    """dist: (L, N, S) importance distributions of L layers over N samples -> (L, L) matrix whose
    upper triangle holds the sample-averaged JS divergence and the rest NaN (notebook cell 16).
    fp32 CUDA input (what the scoring kernels produce) goes through one kernel launch
    (splitq_layer_js); anything else through the torch formulas below."""
    L = dist.shape[0]
    if dist.is_cuda and dist.dtype == torch.float32 and dist.dim() == 3:
        import ctypes as C

        from . import _native as N_
        d = dist.contiguous()
        _, Ns, S = d.shape
        js = torch.full((L, L, Ns), float("nan"), dtype=torch.float64, device=d.device)
        with torch.cuda.device(d.device):
            N_.check(N_.load().splitq_layer_js(C.c_void_p(d.data_ptr()), L, Ns, S, C.c_void_p(js.data_ptr()),
                                               C.c_void_p(torch.cuda.current_stream(d.device).cuda_stream)))
        return js.mean(dim=-1)
    out = torch.full((L, L), float("nan"), dtype=torch.float64, device=dist.device)
    d = dist.double()
    for i in range(L):
        for j in range(i + 1, L):
            out[i, j] = jensen_shannon(d[i], d[j]).mean()
    return out


def _ranks(x: torch.Tensor) -> torch.Tensor:
    order = torch.argsort(x, dim=-1, stable=True)
    ranks = torch.empty_like(order)
    ar = torch.arange(x.shape[-1], device=x.device).expand_as(order)
    ranks.scatter_(-1, order, ar)
    return ranks.double()


def spearman(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """Spearman rank correlation along the last dim (ties broken by index)."""
    ra, rb = _ranks(a), _ranks(b)
    ra = ra - ra.mean(dim=-1, keepdim=True)
    rb = rb - rb.mean(dim=-1, keepdim=True)
    return (ra * rb).sum(dim=-1) / torch.sqrt((ra * ra).sum(dim=-1) * (rb * rb).sum(dim=-1))
