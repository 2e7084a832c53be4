#!/bin/bash
# ncu --set full of the sparse pass (config 4), two launches (levels 3 and 4 of the first timed tree)
mkdir -p gpurun_out
CMD4="python bench.py --steps 6 --warmup 6 --no-cpu --no-e2e --workload c4"
$CMD4 > gpurun_out/plain4.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"eng_hist_csc_kernel" -s 9 -c 2 \
  -o gpurun_out/r02_hist_csc $CMD4 > gpurun_out/ncu_csc.log 2>&1
python scripts/ncu_summary.py gpurun_out/r02_hist_csc.ncu-rep gpurun_out/r02_hist_csc_ncu_full.csv
tail -1 gpurun_out/ncu_csc.log | cut -c1-150; ls -la gpurun_out/*.ncu-rep
