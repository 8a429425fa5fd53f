set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/bench_r1_h_fused.json 2> gpurun_out/bench_h.err; tail -c 600 gpurun_out/bench_h.err
timeout 600 python bench.py --importance materialised > gpurun_out/bench_r1_h_materialised.json 2>> gpurun_out/bench_h.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1_h.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --e2e-steps 1 > gpurun_out/ncu_h.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fused_importance_kernel --launch-skip 4 -c 2 -f -o gpurun_out/k2h_full python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1 > gpurun_out/ncu_h_full.log 2>&1
ls -la gpurun_out | tail -8
