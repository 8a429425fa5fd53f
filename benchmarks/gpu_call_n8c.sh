mkdir -p gpurun_out
nvidia-smi topo -m 2>/dev/null | head -14
lscpu | grep -i "numa\|socket\|^CPU(s)" | head -8
for v in "" "--no-numa-bind"; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 8 --steps 20 --warmup 5 --e2e-steps 10 $v 2>gpurun_out/n8c.err | grep '^{' > gpurun_out/bench_r2_n8c.json
python - <<PY
import json
b=json.loads(open('gpurun_out/bench_r2_n8c.json').read())
print('[$v]', round(b['value']/1e6,1), 'M', round(b['ms_per_step'],4), 'e2e', round(b['e2e']['value']/1e6,2), 'M', b['e2e'].get('host_placement'))
PY
done
tail -3 gpurun_out/n8c.err
