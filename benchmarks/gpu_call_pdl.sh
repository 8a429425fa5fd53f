mkdir -p gpurun_out
for v in pdl nopdl pdl nopdl; do
  if [ $v = pdl ]; then unset SPLITQ_LIB; else export SPLITQ_LIB=$PWD/llm-inference-in-distributed-edge-networks_b200/lib/variants/libsplitq_nopdl.so; fi
  echo "== $v"
  SPLITQ_K2_GROUPS=3 timeout 300 python benchmarks/k2_experiments.py --short 2>&1 | head -1
  timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 2 2>/dev/null | python -c "
import sys,json
d=json.loads([x for x in sys.stdin if x.startswith('{')][-1]); print('bench', d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'])"
done
