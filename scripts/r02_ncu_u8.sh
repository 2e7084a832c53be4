#!/bin/bash
# ncu --set full of the u8 histogram pass (config 3's matrix pre-binned), three launches (levels 1, 4, 7 of a tree)
mkdir -p gpurun_out
CMD="python bench.py --steps 8 --warmup 8 --no-cpu --no-e2e --workload c3-u8"
$CMD > gpurun_out/plain_u8.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"eng_hist_kernel" -s 9 -c 8 \
  -o /tmp/r02_hist_u8 $CMD > gpurun_out/ncu_u8.log 2>&1
python scripts/ncu_summary.py /tmp/r02_hist_u8.ncu-rep gpurun_out/r02_hist_u8_ncu_full.csv
tail -1 gpurun_out/ncu_u8.log | cut -c1-150
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/r02_hist_u8_ncu_full.csv')))
h=rows[0]
for i,k in enumerate(h): print(k,'=',[r[i][:12] for r in rows[2:]])
PY
