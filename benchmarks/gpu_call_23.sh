mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_channel_quant.py tests/test_evalloop.py tests/test_kernels_gpu.py -q -m gpu 2>&1 | tail -12
