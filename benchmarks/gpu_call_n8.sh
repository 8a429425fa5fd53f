mkdir -p gpurun_out
for t in 1 2; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2953$t bench.py --gpus 8 --steps 20 --warmup 5 2>gpurun_out/n8_$t.err | grep '^{' > gpurun_out/bench_r2_n8_$t.json
python - <<PY
import json
b=json.loads(open('gpurun_out/bench_r2_n8_$t.json').read())
print('run $t', {k:b[k] for k in ('value','ms_per_step')}, b['roofline']['kernel_ms'], 'e2e', b['e2e']['value'], b['median_step'])
for r in b['per_rank_ms']: print(' ', r['rank'], 'step', {k:round(v,3) for k,v in r['step_ms'].items()}, 'score med/max', round(r['score_ms']['median'],3), round(r['score_ms']['max'],3), 'gap max', round(r['main_gap_ms']['max'],3), 'wait max', round(r['partner_wait_ms']['max'],3), 'host max', round(r['host_submit_ms']['max'],3), 'ahead', round(r['host_submit_ms']['host_ran_ahead_ms'],2))
PY
done
