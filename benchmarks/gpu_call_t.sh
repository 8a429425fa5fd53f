mkdir -p gpurun_out
timeout 600 python benchmarks/k2_margin.py 2>&1 | tail -8
timeout 600 python -m pytest tests/test_fused_gpu.py -q -x -m gpu -k general_epilogue_forced 2>&1 | tail -3
