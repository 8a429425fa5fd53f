mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none -k regex:quant_pack_ring --launch-skip 3 -c 1 -f -o gpurun_out/qp_ring python benchmarks/side_kernels.py > gpurun_out/qp_ring.log 2>&1
timeout 120 ncu -i gpurun_out/qp_ring.ncu-rep --page raw --csv > gpurun_out/qp_ring_raw.csv 2>/dev/null
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/qp_ring_raw.csv')))
hdr=rows[0]; u=rows[1]; r=rows[2]
keys=['gpu__time_duration.sum','dram__throughput.avg.pct','gpu__dram_throughput.avg.pct','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct','launch__occupancy_limit','launch__waves','achieved_occupancy','launch__grid_size','launch__block_size','smsp__inst_executed.sum','l1tex__throughput.avg.pct','lts__throughput.avg.pct','launch__registers','shared_mem','smsp__pcsamp_warps_issue_stalled']
for i,h in enumerate(hdr):
    if any(k in h for k in keys): print(h, r[i], u[i])
PY
