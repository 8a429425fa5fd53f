mkdir -p gpurun_out
for v in base sel512 sel256 base sel512; do
  if [ $v = base ]; then unset SPLITQ_LIB; else export SPLITQ_LIB=$PWD/llm-inference-in-distributed-edge-networks_b200/lib/variants/libsplitq_$v.so; fi
  echo "== $v"
  timeout 300 python -m pytest tests/test_kernels_gpu.py -q -x -m gpu -k "select" 2>&1 | tail -1
  timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:select_bits -c 12 --csv --log-file gpurun_out/sel.csv python benchmarks/side_kernels.py > /dev/null 2>&1
  grep select_bits gpurun_out/sel.csv | awk -F'","' '{print $NF}' | tr '\n' ' '; echo
done
