mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_kernels_gpu.py tests/test_zz_no_alloc.py tests/test_cuda_aten_parity_gpu.py -q -x -m gpu 2>&1 | tail -2
SPLITQ_QP_RING=1 timeout 1500 python -m pytest tests/test_kernels_gpu.py -q -x -m gpu 2>&1 | tail -2
for r in 0 1 0 1; do
  if [ $r = 1 ]; then export SPLITQ_QP_RING=1; else unset SPLITQ_QP_RING; fi
  echo "== ring $r"; timeout 300 python benchmarks/side_kernels.py 2>&1 | grep -E "^quant_pack"
  timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:quant_pack -c 6 --csv --log-file gpurun_out/qp.csv python benchmarks/side_kernels.py > /dev/null 2>&1
  grep quant_pack gpurun_out/qp.csv | awk -F'","' '{print $NF}' | tr '\n' ' '; echo
done
