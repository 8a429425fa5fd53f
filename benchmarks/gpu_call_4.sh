mkdir -p gpurun_out
for v in base p2 p3 p4 p5 p6 p4g3; do
  echo "== $v"
  if [ $v = base ]; then unset SPLITQ_LIB; else export SPLITQ_LIB=$PWD/llm-inference-in-distributed-edge-networks_b200/lib/variants/libsplitq_$v.so; fi
  timeout 200 python benchmarks/k2_experiments.py --short 2>&1 | tail -6
done > gpurun_out/k2_variants.txt 2>&1
cat gpurun_out/k2_variants.txt
export SPLITQ_LIB=$PWD/llm-inference-in-distributed-edge-networks_b200/lib/variants/libsplitq_p4.so
timeout 600 python -m pytest tests/test_fused_gpu.py -q 2>&1 | tail -4
