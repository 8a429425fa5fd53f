mkdir -p gpurun_out
timeout 900 python benchmarks/k2_stress.py 60 2>&1 | grep -E "MISMATCH|TOTAL"
for i in 1 2 3 4; do timeout 600 python -m pytest tests/test_fused_gpu.py -q -x -m gpu -p no:randomly 2>&1 | tail -25 | grep -E "passed|failed|assert|Error|error" | head -8; done
