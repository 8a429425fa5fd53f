"""K2 alone over the microbenchmark grid's shapes: median ms per call, CUDA events, cold L2 by
rotating operand sets.  SPLITQ_LIB picks the library build."""
import os, sys, statistics, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "llm-inference-in-distributed-edge-networks_b200"))
import splitq

def run(B, H, KVH, S, causal=True, reps=12):
    g = torch.Generator(device="cuda").manual_seed(0)
    sets = [(torch.randn(B, H, S, 64, device="cuda", generator=g), torch.randn(B, KVH, S, 64, device="cuda", generator=g)) for _ in range(3)]
    for q, k in sets:
        splitq.importance_fused(q, k, 0.125, causal=causal)
    torch.cuda.synchronize()
    ts = []
    for i in range(reps):
        q, k = sets[i % 3]
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); splitq.importance_fused(q, k, 0.125, causal=causal); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return statistics.median(ts)

if __name__ == "__main__":
    tokens = 32768
    for (H, KVH) in ((14, 2), (8, 8)):
        for S in (128, 256, 512, 1024, 2048, 4096, 8192, 16384):
            B = max(1, tokens // S)
            print(f"H={H:2d} KVH={KVH} B={B:3d} S={S:5d}  {run(B, H, KVH, S):.4f} ms", flush=True)
    print(f"H=14 KVH=2 B=  1 S= 4096  {run(1, 14, 2, 4096):.4f} ms  (one window)")
    print(f"H= 8 KVH=8 B=  1 S= 2048  {run(1, 8, 8, 2048):.4f} ms  (one window)")
