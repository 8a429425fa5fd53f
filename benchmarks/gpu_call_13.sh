mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_kernels_gpu.py -q -x -k "staged or full_size or fixtures" 2>&1 | tail -3
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/side_launches.csv python benchmarks/side_kernels.py > gpurun_out/side_ncu.log 2>&1
grep "quant_pack" gpurun_out/side_launches.csv | tail -3 | cut -d, -f5,13-
