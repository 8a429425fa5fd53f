mkdir -p gpurun_out
timeout 600 python benchmarks/window_latency.py > gpurun_out/r2_window_latency.json 2> gpurun_out/wl.err; tail -2 gpurun_out/wl.err; head -c 900 gpurun_out/r2_window_latency.json; echo
for w in pythia-70m_split3_lastrow_s2048_int4 pythia-70m_split3_colmean_s512_int4group; do
timeout 600 python bench.py --workload $w --no-cpu-baseline 2>/dev/null | grep "^{" > gpurun_out/bench_$w.json; python -c "
import json,sys
d=json.loads(open('gpurun_out/bench_$w.json').read()); print('$w', round(d['value']/1e6,1), 'M', round(d['ms_per_step'],4))"
done
