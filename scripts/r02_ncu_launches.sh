#!/bin/bash
# launch list of the default bench command (every kernel of the timed region with its duration and DRAM bytes)
mkdir -p gpurun_out
CMD="python bench.py --steps 8 --warmup 8 --no-cpu --no-e2e"
$CMD > gpurun_out/plain.log 2> gpurun_out/plain.err && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 150 -c 260 --csv \
  --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
python scripts/launch_summary.py gpurun_out/r02_launches.csv gpurun_out/r02_launch_summary.csv
cat gpurun_out/r02_launch_summary.csv | cut -c1-150 | head -12
