#!/bin/bash
# ncu --set full of the dense histogram pass over one tree (levels 1..7 and the next tree's level 0)
mkdir -p gpurun_out
CMD="python bench.py --steps 8 --warmup 8 --no-cpu --no-e2e"
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"eng_hist_kernel" -s 9 -c 8 \
  -o /tmp/r02_eng_hist $CMD > gpurun_out/ncu_hist.log 2>&1
python scripts/ncu_summary.py /tmp/r02_eng_hist.ncu-rep gpurun_out/r02_eng_hist_ncu_full.csv
ncu -i /tmp/r02_eng_hist.ncu-rep --page details --csv 2>/dev/null | grep -E "Estimated Speedup|Rule" | head -60 > gpurun_out/r02_eng_hist_rules.txt
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/r02_eng_hist_ncu_full.csv')))
h=rows[0]
for i,k in enumerate(h): print(k,'=',[r[i][:10] for r in rows[2:]])
PY
