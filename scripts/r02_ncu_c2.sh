#!/bin/bash
# ncu --set full of config 2's histogram pass (1M x 28): six launches (one tree)
mkdir -p gpurun_out
CMD="python bench.py --steps 12 --warmup 6 --no-cpu --no-e2e --workload c2"
$CMD > gpurun_out/plain_c2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"eng_hist_kernel" -s 7 -c 6 \
  -o /tmp/r02_hist_c2 $CMD > gpurun_out/ncu_c2.log 2>&1
python scripts/ncu_summary.py /tmp/r02_hist_c2.ncu-rep gpurun_out/r02_hist_c2_ncu_full.csv
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/r02_hist_c2_ncu_full.csv')))
h=rows[0]
for i,k in enumerate(h): print(k,'=',[r[i][:10] for r in rows[2:]])
PY
ncu -i /tmp/r02_hist_c2.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin))
h=rows[0]
want=[i for i,k in enumerate(h) if 'stall' in k.lower() and 'pct' not in k.lower() and 'ratio' in k.lower()]
for i in want: print(h[i], [r[i][:6] for r in rows[2:5]])
" | sort -t"'" -k2 -r | head -24
