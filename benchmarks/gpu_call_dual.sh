mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_fused_gpu.py tests/test_hf_hook_gpu.py -q -x -m gpu > gpurun_out/dual_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/dual_tests.log
tail -4 gpurun_out/dual_tests.log
timeout 600 python benchmarks/k2_shapes.py > gpurun_out/dual_shapes.log 2>&1; cat gpurun_out/dual_shapes.log
timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 2 > gpurun_out/dual_bench.json 2> gpurun_out/dual_bench.err
python - <<'PY'
import json
d=json.loads([x for x in open('gpurun_out/dual_bench.json') if x.startswith('{')][-1])
print(d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['gpu_launches'])
PY
