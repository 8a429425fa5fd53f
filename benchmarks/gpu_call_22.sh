mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fused_gpu.py tests/test_hf_hook_gpu.py tests/test_kernels_gpu.py -q -x 2>&1 | tail -3
for i in 1 2; do
timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 2 2>/dev/null | grep '^{' > gpurun_out/b_x.json
python - <<'PY'
import json
b=json.loads(open('gpurun_out/b_x.json').read())
print(round(b['value']/1e6,2), 'M', round(b['ms_per_step'],4), 'K2', round(b['roofline']['kernel_ms'],4), b['gpu_launches'])
PY
done
