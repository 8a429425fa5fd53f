mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_hf_hook_gpu.py tests/test_fused_gpu.py -q 2>&1 | tail -3
for fam in qwen pythia; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --eval-family $fam --steps 6 --warmup 1 2>gpurun_out/eval_${fam}_n2.err | grep '^{' > gpurun_out/eval_${fam}_n2.json; cut -c1-330 gpurun_out/eval_${fam}_n2.json; tail -2 gpurun_out/eval_${fam}_n2.err
timeout 900 python bench.py --eval-family $fam --steps 6 --warmup 1 2>/dev/null | grep '^{' > gpurun_out/eval_${fam}_n1.json; cut -c1-200 gpurun_out/eval_${fam}_n1.json
done
