import sys, torch
import os; sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "llm-inference-in-distributed-edge-networks_b200"))
import splitq
B,H,KVH,S=8,14,2,4096
SHORT = "--short" in sys.argv
if SHORT: sys.argv.remove("--short")
if len(sys.argv) > 2: B,S=int(sys.argv[1]),int(sys.argv[2])
g=torch.Generator(device="cuda").manual_seed(0)
q=torch.randn(B,H,S,64,device="cuda",generator=g); k=torch.randn(B,KVH,S,64,device="cuda",generator=g)
def t(mode,n=10):
    for _ in range(3): splitq.importance_fused(q,k,0.125,mode=mode)
    torch.cuda.synchronize()
    a=torch.cuda.Event(enable_timing=True); b=torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): splitq.importance_fused(q,k,0.125,mode=mode)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b)/n
MODES = [("x3",0),("x3 general epilogue",512),("x3 general nomma",512+32),("x1",1),("x3 nomath",16),("x1 nomath",17),("x3 nomma",32),("x1 nomma",33),("x3 nomath nomma",48),("x1 nomath nomma",49),("x3 nomma noload",96),("x3 noload",64),("x3 no operand fetch",256),("x3 nomath nomma no operand fetch",256+48)]
if SHORT: MODES = [m for m in MODES if m[0] in ("x3", "x3 general epilogue", "x3 nomath", "x3 nomma", "x3 nomath nomma")]
for name,mode in MODES:
    print(f"{name:18s} {t(mode):.3f} ms")

if len(sys.argv) > 3:  # cycle accounting printed by the kernel (debug instantiation)
    splitq.importance_fused(q,k,0.125,mode=int(sys.argv[3])); torch.cuda.synchronize()
