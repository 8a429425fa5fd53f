mkdir -p gpurun_out
i=0
for v in "--depth 2" "--depth 2 --main-priority 1" "--depth 2 --transfer nccl" "--depth 2 --pack-on-main" "--depth 3 --main-priority 1"; do
i=$((i+1))
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2952$i bench.py --gpus 2 --steps 20 --warmup 5 $v --no-cpu-baseline --e2e-steps 2 2>/dev/null | grep '^{' > gpurun_out/bench_r2_n2_v$i.json
python - <<PY
import json
b=json.loads(open('gpurun_out/bench_r2_n2_v$i.json').read())
print('$v', {k:b[k] for k in ('value','ms_per_step')}, b['roofline']['kernel_ms'])
for r in b['per_rank_ms']: print(' ', r['rank'], r['step_ms']['median'], r['score_ms']['median'], 'gap', r['main_gap_ms']['median'], 'wait', r['partner_wait_ms']['median'])
PY
done
