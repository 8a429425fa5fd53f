mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu 2>&1 | tail -5
timeout 300 python benchmarks/side_kernels.py 2>&1 | head -6
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
