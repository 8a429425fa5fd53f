mkdir -p gpurun_out
for v in base d4m10 d4m8 d8m5 d4m12; do
  if [ $v = base ]; then unset SPLITQ_LIB; else export SPLITQ_LIB=$PWD/llm-inference-in-distributed-edge-networks_b200/lib/variants/libsplitq_$v.so; fi
  timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/sl_$v.csv python benchmarks/side_kernels.py > /dev/null 2>&1
  echo "$v qp $(grep quant_pack_kernel gpurun_out/sl_$v.csv | tail -3 | awk -F'","' '{print $NF}' | tr -d '"' | sort -n | head -1) dq $(grep 'dequant_kernel' gpurun_out/sl_$v.csv | tail -5 | awk -F'","' '{print $NF}' | tr -d '"' | sort -n | head -3 | tr '\n' ' ')"
done
