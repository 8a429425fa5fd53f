mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_kernels_gpu.py tests/test_zz_no_alloc.py tests/test_cuda_aten_parity_gpu.py tests/test_channel_quant.py tests/test_models.py -q -x -m gpu 2>&1 | tail -3
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"dequant_kernel|quant_pack_kernel" -c 12 --csv --log-file gpurun_out/dq.csv python benchmarks/side_kernels.py > /dev/null 2>&1
grep -E "dequant|quant_pack" gpurun_out/dq.csv | awk -F'","' '{print $5, $NF}' | cut -c1-40,100- | tail -12
timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 2 > gpurun_out/dq_bench.json 2>gpurun_out/dq_bench.err; tail -3 gpurun_out/dq_bench.err
python - <<'PY'
import json
d=json.loads([x for x in open('gpurun_out/dq_bench.json') if x.startswith('{')][-1])
print('bench', d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'])
for k,v in d['roofline_aux'].items():
    if isinstance(v,dict): print('   ',k, round(v['kernel_us'],2), round(v['frac'],3))
    else: print('   ',k,v)
PY
