mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_fused_gpu.py -q -x -m gpu -k "mixed_items or alignments or prescaled or large_norms" 2>&1 | tail -15
