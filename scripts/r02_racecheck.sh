#!/bin/bash
mkdir -p gpurun_out
python scripts/racecheck_small.py > gpurun_out/racecheck_plain.log 2>&1 && timeout 1500 compute-sanitizer --tool racecheck --racecheck-report all --print-limit 20 python scripts/racecheck_small.py > gpurun_out/r02_racecheck.txt 2>&1
tail -15 gpurun_out/r02_racecheck.txt | cut -c1-200
