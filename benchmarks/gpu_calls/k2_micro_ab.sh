mkdir -p gpurun_out
for v in base eb32 eb8 p2 st5 base; do
  if [ $v = base ]; then unset SPLITQ_LIB; else export SPLITQ_LIB=$PWD/llm-inference-in-distributed-edge-networks_b200/lib/variants/libsplitq_$v.so; fi
  echo "== $v"
  SPLITQ_K2_GROUPS=3 timeout 300 python benchmarks/k2_experiments.py --short 2>&1 | head -1
  timeout 300 python bench.py --steps 40 --warmup 5 --no-cpu-baseline --e2e-steps 0 2>/dev/null | python -c "
import sys,json
d=json.loads([x for x in sys.stdin if x.startswith('{')][-1]); print('bench', round(d['value']/1e6,2), round(d['ms_per_step'],4), round(d['roofline']['kernel_ms'],4))"
done
