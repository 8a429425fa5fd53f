mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fused_gpu.py tests/test_hf_hook_gpu.py -q -x 2>&1 | tail -15
timeout 300 python benchmarks/k2_experiments.py > gpurun_out/k2_fast.txt 2>&1; cat gpurun_out/k2_fast.txt
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_r2_fast.json 2> gpurun_out/bench_r2_fast.err; tail -c 300 gpurun_out/bench_r2_fast.err
python - <<'PY'
import json
b=json.load(open('gpurun_out/bench_r2_fast.json'))
print({k:b[k] for k in ('value','ms_per_step')}, b['roofline']['kernel_ms'], b['clocks'])
PY
