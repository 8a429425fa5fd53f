mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/final_tests.log 2>&1; tail -4 gpurun_out/final_tests.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
timeout 600 python bench.py 2>/dev/null | grep "^{" > gpurun_out/final_bench.json; python -c "
import json
d=json.loads(open('gpurun_out/final_bench.json').read()); print(round(d['value']/1e6,2),'M', round(d['ms_per_step'],4), 'k2', round(d['roofline']['kernel_ms'],4), round(d['roofline']['frac'],3), 'e2e', round(d['e2e']['value']/1e6,2), 'cpu', round(d['cpu_baseline']['value']))"
