#!/bin/bash
# GPU-box helper (development): engine histogram kernel variants on config 3, hist ms per level.
mkdir -p gpurun_out
for v in "$@"; do
  H2OK_ENG_VARIANT=$v timeout 300 python bench.py --path engine --no-cpu --no-e2e --steps 8 --warmup 8 > gpurun_out/v$v.json 2>gpurun_out/v$v.err || tail -3 gpurun_out/v$v.err
  python - <<PY
import json
d=json.load(open("gpurun_out/v$v.json"))
print("variant $v: ms/step %.2f frac %.3f" % (d["ms_per_step"], d["roofline"]["frac"]), " ".join("%.2f" % x["hist_ms"] for x in d["config"]["per_level"].values()))
PY
done
