mkdir -p gpurun_out
bash benchmarks/gpu_call_9.sh 2>&1 | tail -4
unset SPLITQ_LIB
timeout 1200 python -m pytest tests/test_selection_analytics.py tests/test_kernels_gpu.py tests/test_property_gpu.py -q 2>&1 | tail -5
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/side_launches.csv python benchmarks/side_kernels.py > gpurun_out/side_ncu.log 2>&1
grep select_bits gpurun_out/side_launches.csv | tail -3
