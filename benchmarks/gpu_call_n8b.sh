mkdir -p gpurun_out
for fam in qwen pythia; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 8 --eval-family $fam --steps 6 --warmup 1 2>gpurun_out/eval_${fam}_n8.err | grep '^{' > gpurun_out/eval_${fam}_n8.json; cut -c1-200 gpurun_out/eval_${fam}_n8.json
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus 4 --steps 20 --warmup 5 2>/dev/null | grep '^{' > gpurun_out/bench_r2_n4.json; cut -c1-160 gpurun_out/bench_r2_n4.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29553 bench.py --gpus 8 --steps 20 --warmup 5 2>/dev/null | grep '^{' > gpurun_out/bench_r2_n8_3.json; cut -c1-160 gpurun_out/bench_r2_n8_3.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29554 benchmarks/layer_sweep.py --family qwen --windows 32 > gpurun_out/r2_layer_sweep_n8.json 2>gpurun_out/ls.err; tail -c 300 gpurun_out/r2_layer_sweep_n8.json; tail -2 gpurun_out/ls.err
