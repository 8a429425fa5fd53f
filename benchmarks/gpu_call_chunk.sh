mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fused_gpu.py tests/test_hf_hook_gpu.py -q -x -m gpu > gpurun_out/chunk_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/chunk_tests.log
tail -3 gpurun_out/chunk_tests.log
timeout 300 python benchmarks/k2_experiments.py --short > gpurun_out/chunk_k2.log 2>&1
SPLITQ_K2_NO_CHUNKS=1 timeout 300 python benchmarks/k2_experiments.py --short > gpurun_out/chunk_k2_off.log 2>&1
timeout 300 python benchmarks/k2_experiments.py --short > gpurun_out/chunk_k2_b.log 2>&1
tail -6 gpurun_out/chunk_k2.log; tail -6 gpurun_out/chunk_k2_off.log; tail -6 gpurun_out/chunk_k2_b.log
timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 2 > gpurun_out/chunk_bench.json 2> gpurun_out/chunk_bench.err
SPLITQ_K2_NO_CHUNKS=1 timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 2 > gpurun_out/chunk_bench_off.json 2>> gpurun_out/chunk_bench.err
tail -c 1500 gpurun_out/chunk_bench.json; tail -c 1500 gpurun_out/chunk_bench_off.json
