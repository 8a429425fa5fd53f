mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_selection_analytics.py tests/test_kernels_gpu.py tests/test_property_gpu.py tests/test_zz_no_alloc.py -q 2>&1 | tail -12
timeout 300 python benchmarks/side_kernels.py > gpurun_out/side_kernels.txt 2>&1; cat gpurun_out/side_kernels.txt | head -8
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/side_launches.csv python benchmarks/side_kernels.py > gpurun_out/side_ncu.log 2>&1
python - <<'PY'
import csv, collections
rows=list(csv.reader(open('gpurun_out/side_launches.csv')))
i0=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]
hdr=rows[i0]; d=collections.defaultdict(lambda: collections.defaultdict(list))
for r in rows[i0+1:]:
    if len(r)<len(hdr): continue
    d[r[hdr.index('Kernel Name')][:60]][r[hdr.index('Metric Name')]].append(float(r[hdr.index('Metric Value')].replace(',','')))
for k,v in d.items():
    if 'zero' in k.lower() or 'fill' in k.lower(): continue
    print(k, {m:(sorted(x)[len(x)//2]) for m,x in v.items()}, hdr[hdr.index('Metric Unit')] if 'Metric Unit' in hdr else '')
PY
