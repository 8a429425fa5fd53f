mkdir -p gpurun_out
for v in base ff1; do if [ $v = base ]; then unset SPLITQ_LIB; else export SPLITQ_LIB=$PWD/llm-inference-in-distributed-edge-networks_b200/lib/variants/libsplitq_ff1.so; fi; echo "== $v"
K2_ONE_REPS=8 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/fuse_launches.csv python benchmarks/k2_one.py > /dev/null 2>&1
python - <<'PY'
import csv,collections,sys
rows=list(csv.reader(open('gpurun_out/fuse_launches.csv')))
for i,r in enumerate(rows):
    if 'Kernel Name' in r: hdr=r; start=i+1; break
kn=hdr.index('Kernel Name'); mv=hdr.index('Metric Value')
agg=collections.OrderedDict()
for r in rows[start:]:
    if len(r)<=mv: continue
    agg.setdefault(r[kn].split('(')[0][:60],[]).append(float(r[mv].replace(',','')))
for n,v in agg.items(): print(f"{n:62s} n={len(v):3d} median={sorted(v)[len(v)//2]/1e3:9.2f} us")
PY
done
unset SPLITQ_LIB
for m in ; do SPLITQ_K2_GROUPS=3 timeout 100 python - $m <<'PY'
import sys, os, torch, statistics
sys.path.insert(0, "llm-inference-in-distributed-edge-networks_b200")
import splitq
from splitq import _native as N
dbg=int(sys.argv[1])
B,H,KVH,S=8,14,2,4096
g=torch.Generator(device="cuda").manual_seed(0)
q=torch.randn(B,H,S,64,device="cuda",generator=g); k=torch.randn(B,KVH,S,64,device="cuda",generator=g)
mode=N.FUSED_F16X3|(dbg<<4)
for _ in range(3): splitq.importance_fused(q,k,0.125,mode=mode)
torch.cuda.synchronize(); ts=[]
for _ in range(15):
    a,b=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    a.record(); splitq.importance_fused(q,k,0.125,mode=mode); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
print("dbg",dbg,"%.4f ms"%statistics.median(ts))
PY
done
