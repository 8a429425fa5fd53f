# what the driver runs at round end, plus the evidence files kept under profiles/ (run on a B200 box)
set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/bench_fused.json 2> gpurun_out/bench.err; tail -c 600 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_reference.json 2>> gpurun_out/bench.err
timeout 600 python bench.py --importance materialised > gpurun_out/bench_materialised.json 2>> gpurun_out/bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_fused.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1 > gpurun_out/ncu.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_importance_kernel --launch-skip 6 -c 2 -f -o gpurun_out/k2_full python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1 > gpurun_out/ncu_full.log 2>&1
timeout 300 ncu -i gpurun_out/k2_full.ncu-rep --page raw --csv > gpurun_out/k2_full_raw.csv 2>/dev/null
timeout 900 python benchmarks/sweep.py --out gpurun_out/sweep.json --md gpurun_out/sweep.md > gpurun_out/sweep.log 2>&1; tail -3 gpurun_out/sweep.log
ls -la gpurun_out | tail -12
