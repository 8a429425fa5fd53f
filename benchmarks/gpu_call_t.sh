mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/full_tests.log 2>&1; tail -60 gpurun_out/full_tests.log
