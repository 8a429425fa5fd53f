mkdir -p gpurun_out
for v in base r112 r104 r96 base; do
  if [ $v = base ]; then unset SPLITQ_LIB; else export SPLITQ_LIB=$PWD/llm-inference-in-distributed-edge-networks_b200/lib/variants/libsplitq_$v.so; fi
  timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 2 2>/dev/null | grep '^{' > gpurun_out/b_$v.json
  python - <<PY
import json
b=json.loads(open('gpurun_out/b_$v.json').read()); r=b['per_rank_ms'][0]
print('$v', round(b['value']/1e6,2), 'M', round(b['ms_per_step'],4), 'K2', round(b['roofline']['kernel_ms'],4), 'step med', round(r['step_ms']['median'],4))
PY
  timeout 100 python benchmarks/k2_experiments.py --short 2>&1 | head -1
done
