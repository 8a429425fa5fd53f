mkdir -p gpurun_out
V=llm-inference-in-distributed-edge-networks_b200/lib/variants
: > gpurun_out/var_k2.log
for n in base g3 p3 g3p3 g3p4 base; do
  if [ $n = base ]; then unset SPLITQ_LIB; else export SPLITQ_LIB=$PWD/$V/libsplitq_$n.so; fi
  echo "== $n" >> gpurun_out/var_k2.log
  timeout 300 python benchmarks/k2_experiments.py --short >> gpurun_out/var_k2.log 2>&1
  timeout 300 python bench.py --steps 40 --warmup 5 --no-cpu-baseline --e2e-steps 0 2>/dev/null | python -c "
import sys,json
l=[x for x in sys.stdin if x.startswith('{')][-1]; d=json.loads(l)
print('bench', d['value'], d['ms_per_step'], d['roofline']['kernel_ms'])" >> gpurun_out/var_k2.log 2>&1
done
cat gpurun_out/var_k2.log
