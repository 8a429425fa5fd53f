mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu 2>&1 | tail -12
timeout 600 python bench.py > gpurun_out/bench_r2_b.json 2> gpurun_out/bench_r2_b.err; tail -c 300 gpurun_out/bench_r2_b.err
python - <<'PY'
import json
b=json.load(open('gpurun_out/bench_r2_b.json'))
print({k:b[k] for k in ('value','ms_per_step')}, b['roofline']['kernel_ms'], b['roofline']['frac'], b['clocks'], b['e2e']['value'], b['cpu_baseline']['value'])
PY
