mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_selection_analytics.py tests/test_kernels_gpu.py tests/test_property_gpu.py tests/test_zz_no_alloc.py -q 2>&1 | tail -12
timeout 300 python benchmarks/side_kernels.py > gpurun_out/side_kernels.txt 2>&1; cat gpurun_out/side_kernels.txt | head -8
