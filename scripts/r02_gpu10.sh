#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_loop.py -m gpu -x -q -k "sparse" 2>&1 | tail -3
for cfg in "4 112" "4 56" "2 56"; do
set -- $cfg
H2OK_CSC_U=$1 H2OK_CSC_KB=$2 timeout 600 python bench.py --steps 12 --warmup 10 --no-cpu --no-e2e --workload c4 > gpurun_out/bench_x.json 2> gpurun_out/bench_x.err
python -c "
import json; d=json.load(open('gpurun_out/bench_x.json')); print('U=$1 KB=$2 ms/step %.3f ' % d['ms_per_step'], ' '.join('L%s:%.2f' % (k, v['hist_ms']) for k,v in d['config']['per_level'].items()))"
done
