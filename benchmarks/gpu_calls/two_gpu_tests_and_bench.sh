mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 20 --warmup 5 2>gpurun_out/n2.err | grep '^{' > gpurun_out/bench_r2i_n2.json
python - <<'PY'
import json
b=json.loads(open('gpurun_out/bench_r2i_n2.json').read())
print('N=2', round(b['value']/1e6,1), 'M', round(b['ms_per_step'],4), 'e2e', round(b['e2e']['value']/1e6,2), 'M', 'k2', b['roofline']['kernel_ms'], b['roofline']['frac'])
PY
tail -2 gpurun_out/n2.err
