mkdir -p gpurun_out
i=0
for v in "--main-priority -3" "--main-priority 1" "--main-priority 1 --pack-on-main" "--main-priority 1 --depth 3" "--main-priority -1"; do
i=$((i+1))
timeout 600 python bench.py --steps 50 --warmup 5 $v --no-cpu-baseline --e2e-steps 2 2>/dev/null | grep '^{' > gpurun_out/bench_r2_n1_v$i.json
python - <<PY
import json
b=json.loads(open('gpurun_out/bench_r2_n1_v$i.json').read())
r=b['per_rank_ms'][0]
print('$v', {k:round(b[k],4) for k in ('value','ms_per_step')}, round(b['roofline']['kernel_ms'],4), 'gap', round(r['main_gap_ms']['median'],4), 'step med', round(r['step_ms']['median'],4))
PY
done
