mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 8 --steps 20 --warmup 5 2>gpurun_out/n8d.err | grep '^{' > gpurun_out/bench_r2i_n8.json
python - <<'PY'
import json
b=json.loads(open('gpurun_out/bench_r2i_n8.json').read())
print('N=8', round(b['value']/1e6,1), 'M', round(b['ms_per_step'],4), 'e2e', round(b['e2e']['value']/1e6,2), 'M', 'k2', b['roofline']['kernel_ms'], round(b['roofline']['frac'],3))
print([ (r['rank'], round(r['total_ms'],2), round(r['step_ms']['median'],4)) for r in b['per_rank_ms']])
PY
tail -2 gpurun_out/n8d.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29562 bench.py --gpus 4 --steps 20 --warmup 5 2>>gpurun_out/n8d.err | grep '^{' > gpurun_out/bench_r2i_n4.json
python - <<'PY'
import json
b=json.loads(open('gpurun_out/bench_r2i_n4.json').read())
print('N=4', round(b['value']/1e6,1), 'M', round(b['ms_per_step'],4), 'e2e', round(b['e2e']['value']/1e6,2), 'M', 'k2', b['roofline']['kernel_ms'], round(b['roofline']['frac'],3))
PY
