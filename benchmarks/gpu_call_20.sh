mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fused_gpu.py tests/test_hf_hook_gpu.py -q -x 2>&1 | tail -12
timeout 200 python benchmarks/k2_experiments.py --short 2>&1 | tail -6
export SPLITQ_LIB=$PWD/llm-inference-in-distributed-edge-networks_b200/lib/variants/libsplitq_n128.so
echo "== n128"
timeout 900 python -m pytest tests/test_fused_gpu.py -q -x 2>&1 | tail -3
timeout 200 python benchmarks/k2_experiments.py --short 2>&1 | tail -6
