import os, sys, torch
sys.path.insert(0, "llm-inference-in-distributed-edge-networks_b200")
import splitq
B, S, D = 8, 4096, 896
g = torch.Generator(device="cuda").manual_seed(0)
imp = torch.softmax(torch.randn(B, S, device="cuda", generator=g) * 2, -1)
ladder = splitq.ladder_from_fractions([(0.25, 2), (0.25, 4), (0.25, 8)], S)
for _ in range(4):
    splitq.select_bits(imp, ladder, D)
    torch.cuda.synchronize()
