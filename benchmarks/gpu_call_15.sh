mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu 2>&1 | tail -6
timeout 600 python bench.py > gpurun_out/bench_r2_g.json 2> gpurun_out/bench_r2_g.err; tail -c 200 gpurun_out/bench_r2_g.err
python - <<'PY'
import json
b=json.loads([l for l in open('gpurun_out/bench_r2_g.json') if l.startswith('{')][-1])
print({k:b[k] for k in ('value','ms_per_step')}, b['roofline']['kernel_ms'], b['e2e']['value'])
for k,v in b['roofline_aux'].items():
    if isinstance(v,dict): print(k, round(v['kernel_us'],1), round(v['frac'],3))
PY
