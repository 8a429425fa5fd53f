mkdir -p gpurun_out
V=llm-inference-in-distributed-edge-networks_b200/lib/variants
: > gpurun_out/k2_shapes.log
for n in base g3 base g3; do
  if [ $n = base ]; then unset SPLITQ_LIB; else export SPLITQ_LIB=$PWD/$V/libsplitq_$n.so; fi
  echo "== $n" >> gpurun_out/k2_shapes.log
  timeout 600 python benchmarks/k2_shapes.py >> gpurun_out/k2_shapes.log 2>&1
done
cat gpurun_out/k2_shapes.log
