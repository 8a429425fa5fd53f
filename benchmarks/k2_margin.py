import sys, os, torch
sys.path.insert(0, "llm-inference-in-distributed-edge-networks_b200"); sys.path.insert(0, "oracle"); sys.path.insert(0, "tests")
import splitq, splitq_oracle as O
from splitq import _native as N
from test_fused_gpu import make_qk, oracle_scores, fp64_scores, max_rel, GENERAL_PATH
for (B,H,KVH,S,sigma) in [(1, 4, 2, 512, 1.0), (2, 6, 3, 333, 1.2), (1, 14, 2, 1024, 2.0)]:
    gen = torch.Generator().manual_seed(S + H)
    q, k = make_qk(gen, B, H, KVH, S, sigma)
    w32c, w32l = oracle_scores(q, k, True)
    w64c, w64l = fp64_scores(q.cuda(), k.cuda())
    for g in ("2","3"):
        os.environ["SPLITQ_K2_GROUPS"]=g
        cg, lg = splitq.importance_fused(q.cuda(), k.cuda(), 0.125, want_last=True, mode=N.FUSED_F16X3 | GENERAL_PATH)
        cf, lf = splitq.importance_fused(q.cuda(), k.cuda(), 0.125, want_last=True)
        print((B,H,KVH,S,sigma), "groups", g, "general vs fp32cpu %.2e/%.2e  vs fp64 %.2e/%.2e | fast vs fp32cpu %.2e/%.2e vs fp64 %.2e/%.2e | g-vs-f %.2e | fp32cpu vs fp64 %.2e/%.2e" % (
            max_rel(cg.cpu(), w32c), max_rel(lg.cpu(), w32l), max_rel(cg, w64c.float()), max_rel(lg, w64l.float()),
            max_rel(cf.cpu(), w32c), max_rel(lf.cpu(), w32l), max_rel(cf, w64c.float()), max_rel(lf, w64l.float()),
            max_rel(cg, cf), max_rel(w32c.cuda(), w64c.float()), max_rel(w32l.cuda(), w64l.float())))
