#!/usr/bin/env python
"""Small, fast pass over every kernel (vector and generic paths, edge shapes) for
`compute-sanitizer --tool memcheck|racecheck python benchmarks/sanitize_smoke.py`."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "llm-inference-in-distributed-edge-networks_b200"))
import torch  # noqa: E402

import splitq  # noqa: E402
from splitq import _native as N  # noqa: E402

dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(0)
for (B, H, KVH, S, D) in [(1, 4, 2, 200, 896), (2, 2, 2, 77, 20), (1, 1, 1, 1, 5), (1, 8, 8, 384, 512)]:
    A = torch.softmax(torch.randn(B, H, S, S, device=dev, generator=g), dim=-1)
    h = torch.randn(B, S, D, device=dev, generator=g)
    col, last = splitq.importance_from_attention(A, want_last=True)
    splitq.importance_from_attention(A, head_w=[0.1] * H, causal_hint=False)
    q = torch.randn(B, H, S, 64, device=dev, generator=g)
    k = torch.randn(B, KVH, S, 64, device=dev, generator=g)
    fc, fl = splitq.importance_fused(q, k, 0.125, want_last=True)
    splitq.importance_fused(q, k, 0.125, causal=False, mode=N.FUSED_F16X1)
    ladder = [(S // 4, 2), (S // 4, 4), (S // 4, 8)]
    for mode in (N.SCALE_GROUP, N.SCALE_TOKEN):
        nbytes = splitq.payload_bytes(S, D, ladder, mode)
        bits, off = splitq.select_bits(col, ladder, D)
        payload = splitq.quant_pack(h, bits, off, mode, nbytes)
        out = splitq.dequant(payload, S, D)
        gam = torch.ones(D, device=dev)
        splitq.dequant_norm(payload, S, D, N.NORM_LAYERNORM, gam, gam, 1e-5, gam, gam)
        splitq.dequant_norm(payload, S, D, N.NORM_RMSNORM, gam, None, 1e-6)
        if mode == N.SCALE_GROUP:
            splitq.quant_pack_group_noalloc(h, bits, off, torch.empty_like(payload))
    for method in ("channel_8", "channel_4", "channel_1_mean", "channel_1_max"):
        splitq.channel_fake_quant(h, method, want_codes=True)
    torch.cuda.synchronize()
print("sanitize smoke ok")
