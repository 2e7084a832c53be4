#!/bin/bash
# launch list of config 2 (1M x 28): every kernel of a few trees with its duration
mkdir -p gpurun_out
CMD="python bench.py --steps 12 --warmup 6 --no-cpu --no-e2e --workload c2"
$CMD > gpurun_out/plain_c2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 400 --csv \
  --log-file gpurun_out/r02_c2_launches.csv $CMD > gpurun_out/ncu_c2.log 2>&1
python scripts/launch_summary.py gpurun_out/r02_c2_launches.csv gpurun_out/r02_c2_launch_summary.csv
cat gpurun_out/r02_c2_launch_summary.csv | cut -c1-160
tail -2 gpurun_out/plain_c2.log | cut -c1-300
