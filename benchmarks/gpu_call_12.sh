mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_r2_f.json 2> gpurun_out/bench_r2_f.err; tail -c 300 gpurun_out/bench_r2_f.err
python - <<'PY'
import json
b=json.loads([l for l in open('gpurun_out/bench_r2_f.json') if l.startswith('{')][-1])
print({k:b[k] for k in ('value','ms_per_step')}, b['roofline']['kernel_ms'], b['roofline']['frac_of_sustained_peak'], b['roofline']['frac_of_burst_peak'], b['e2e']['value'], b['cpu_baseline']['value'])
for k,v in b['roofline_aux'].items():
    if isinstance(v,dict): print(k, round(v['kernel_us'],1), round(v['frac'],3))
PY
timeout 900 python bench.py --eval-family qwen --steps 4 --warmup 1 2>gpurun_out/eval_n1.err | grep '^{' > gpurun_out/eval_qwen_n1.json; cat gpurun_out/eval_qwen_n1.json | cut -c1-400; tail -3 gpurun_out/eval_n1.err
timeout 900 python bench.py --eval-family pythia --steps 4 --warmup 1 2>gpurun_out/eval_p_n1.err | grep '^{' > gpurun_out/eval_pythia_n1.json; cat gpurun_out/eval_pythia_n1.json | cut -c1-300
