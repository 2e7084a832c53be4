#!/bin/bash
# Round-2 profiler evidence (one GPU): launch list of the default bench command, then ncu --set full of the engine's
# histogram pass, routing and regrouping kernels over one tree (levels 0..7), then the sparse pass of config 4.
# The reports are summarised on the box (scripts/ncu_summary.py) and only the CSVs come back.
mkdir -p gpurun_out
CMD="python bench.py --steps 8 --warmup 8 --no-cpu --no-e2e"
$CMD > gpurun_out/plain.log 2> gpurun_out/plain.err && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 150 -c 260 --csv \
  --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"eng_hist_kernel" -s 9 -c 8 \
  -o /tmp/r02_eng_hist $CMD > gpurun_out/ncu_hist.log 2>&1
python scripts/ncu_summary.py /tmp/r02_eng_hist.ncu-rep gpurun_out/r02_eng_hist_ncu_full.csv
ncu -i /tmp/r02_eng_hist.ncu-rep --page details --csv 2>/dev/null | grep -E "Estimated Speedup|Rule" | head -60 > gpurun_out/r02_eng_hist_rules.txt
$CMD > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"eng_route_kernel|eng_scatter_kernel" -s 17 -c 15 \
  -o /tmp/r02_route_scatter $CMD > gpurun_out/ncu_route.log 2>&1
python scripts/ncu_summary.py /tmp/r02_route_scatter.ncu-rep gpurun_out/r02_route_scatter_ncu_full.csv
CMD4="python bench.py --steps 6 --warmup 6 --no-cpu --no-e2e --workload c4"
$CMD4 > gpurun_out/plain4.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"eng_hist_csc_kernel" -s 6 -c 6 \
  -o /tmp/r02_hist_csc $CMD4 > gpurun_out/ncu_csc.log 2>&1
python scripts/ncu_summary.py /tmp/r02_hist_csc.ncu-rep gpurun_out/r02_hist_csc_ncu_full.csv
tail -1 gpurun_out/ncu_hist.log | cut -c1-150; tail -1 gpurun_out/ncu_route.log | cut -c1-150; tail -1 gpurun_out/ncu_csc.log | cut -c1-150
ls -la gpurun_out | head -30
