mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fused_gpu.py tests/test_hf_hook_gpu.py -q 2>&1 | tail -8
timeout 300 python benchmarks/gap_probe.py > gpurun_out/gap_probe.txt 2>&1; cat gpurun_out/gap_probe.txt
