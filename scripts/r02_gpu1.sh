#!/bin/bash
# Round 2, call 1: baseline bench line of the round-1 engine + ncu --set full of the routing / regroup kernels (100M x 64).
mkdir -p gpurun_out
CMD="python bench.py --steps 8 --warmup 8 --no-cpu --no-e2e"
$CMD > gpurun_out/r02_base_plain.log 2> gpurun_out/r02_base_plain.err && ncu --set full --clock-control none --import-source on \
  -k regex:"eng_route_kernel|eng_scatter_kernel" -s 16 -c 16 -o gpurun_out/r02_route_base $CMD > gpurun_out/ncu_route.log 2>&1
tail -2 gpurun_out/ncu_route.log | cut -c1-200
cat gpurun_out/r02_base_plain.log | cut -c1-400
