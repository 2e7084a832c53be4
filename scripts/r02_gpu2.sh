#!/bin/bash
# parity tests, then the bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu.txt; cat gpurun_out/pytest_gpu.txt
timeout 600 python bench.py --steps 16 --warmup 8 --no-cpu --no-e2e "$@" > gpurun_out/bench_tmp.json 2> gpurun_out/bench_tmp.err; tail -5 gpurun_out/bench_tmp.err
python -c "
import json; d=json.load(open('gpurun_out/bench_tmp.json')); print('value %.1f Gcells/s  ms/step %.3f  frac %.3f' % (d['value']/1e9, d['ms_per_step'], d['roofline']['frac'])); print(' '.join('L%s:%.2f' % (k, v['hist_ms']) for k,v in d['config']['per_level'].items()))"
