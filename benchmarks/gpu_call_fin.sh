mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_fused_gpu.py tests/test_hf_hook_gpu.py -q -x -m gpu 2>&1 | tail -2
for n in old new old new; do
  if [ $n = new ]; then unset SPLITQ_LIB; else export SPLITQ_LIB=$PWD/llm-inference-in-distributed-edge-networks_b200/lib/variants/libsplitq_old.so; fi
  K2_ONE_REPS=8 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/fin_launches_$n.csv python benchmarks/k2_one.py > /dev/null 2>&1
  echo "== $n"
  python - $n <<'PY'
import csv,collections,sys
rows=list(csv.reader(open(f'gpurun_out/fin_launches_{sys.argv[1]}.csv')))
for i,r in enumerate(rows):
    if 'Kernel Name' in r: hdr=r; start=i+1; break
kn=hdr.index('Kernel Name'); mv=hdr.index('Metric Value')
agg=collections.OrderedDict()
for r in rows[start:]:
    if len(r)<=mv: continue
    agg.setdefault(r[kn].split('(')[0][:60],[]).append(float(r[mv].replace(',','')))
for n,v in agg.items(): print(f"{n:62s} n={len(v):3d} median={sorted(v)[len(v)//2]/1e3:9.2f} us")
PY
  SPLITQ_K2_GROUPS=3 timeout 300 python benchmarks/k2_experiments.py --short 2>&1 | head -1
done
