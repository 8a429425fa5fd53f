#!/usr/bin/env python
"""Times the side-stream kernels of the bench step alone (B = 8 Qwen2-0.5B windows, S = 4096, D = 896,
8/4/2-bit ladder, per-token scale): select, quantise+pack, device copy, dequantise; prints us and the
fraction of the measured HBM copy bandwidth each reaches on its algorithmic bytes."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "llm-inference-in-distributed-edge-networks_b200"))
import torch
import splitq
from splitq import _native as N

B, S, D = 8, 4096, 896
if len(sys.argv) > 3: B, S, D = (int(x) for x in sys.argv[1:4])
peak = 6554.6
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
h = torch.randn(B, S, D, device=dev, generator=g)
imp = torch.rand(B, S, device=dev, generator=g)
ladder = splitq.ladder_from_fractions([(0.25, 2), (0.25, 4), (0.25, 8)], S)
nbytes = splitq.payload_bytes(S, D, ladder, N.SCALE_TOKEN)
bits, off = splitq.select_bits(imp, ladder, D)
payload = torch.empty(B, nbytes, dtype=torch.uint8, device=dev)
payload2 = torch.empty_like(payload)
out = torch.empty(B, S, D, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def timed(fn, n=30):
    ts = []
    for _ in range(n + 3):
        flush.zero_()                      # cold L2: the working set of one call (160 MB) would otherwise partly stay
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts = sorted(ts[3:])
    return ts[len(ts) // 2]

rows = [
    ("select_bits", lambda: splitq.select_bits(imp, ladder, D, bits, off), 4 * B * S + 5 * B * S),
    ("quant_pack", lambda: splitq.quant_pack(h, bits, off, N.SCALE_TOKEN, nbytes, payload), 4 * B * S * D + 5 * B * S + B * nbytes),
    ("copy (hop stand-in)", lambda: splitq.transfer(payload2, payload), 2 * B * nbytes),
    ("dequant", lambda: splitq.dequant(payload, S, D, out), B * nbytes + 4 * B * S * D),
]
hc = torch.randn(B, S, D, device=dev, generator=g)
outc = torch.empty_like(hc)
rows.append(("channel_fake_quant (channel_4)", lambda: splitq.channel_fake_quant(hc, "channel_4", out=outc), 3 * 4 * B * S * D))
res = {}
for name, fn, alg in rows:
    us = timed(fn)
    res[name] = {"us": us, "algorithmic_bytes": alg, "GBps": alg / us / 1e3, "frac_of_measured_hbm": alg / us / 1e3 / peak}
    print(f"{name:22s} {us:8.2f} us  {alg / 1e6:8.2f} MB  {alg / us / 1e3:8.1f} GB/s  {alg / us / 1e3 / peak:5.2f} of {peak:.0f}")
print(json.dumps(res))
