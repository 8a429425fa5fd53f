mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_fused_gpu.py tests/test_hf_hook_gpu.py -q -x -m gpu 2>&1 | tail -12
for i in 1 2; do SPLITQ_K2_GROUPS=3 timeout 300 python benchmarks/k2_experiments.py --short 2>&1 | head -1; done
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
timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 2 2>/dev/null | python -c "
import sys,json
d=json.loads([x for x in sys.stdin if x.startswith('{')][-1]); print('bench', d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'])"
