import sys, os, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "llm-inference-in-distributed-edge-networks_b200"))
import splitq
B, H, KVH, S = 8, 14, 2, 4096
mode = int(sys.argv[1]) if len(sys.argv) > 1 else 0
g = torch.Generator(device="cuda").manual_seed(0)
q = torch.randn(B, H, S, 64, device="cuda", generator=g); k = torch.randn(B, KVH, S, 64, device="cuda", generator=g)
for _ in range(int(os.environ.get("K2_ONE_REPS", "3"))):
    splitq.importance_fused(q, k, 0.125, mode=mode)
torch.cuda.synchronize()
