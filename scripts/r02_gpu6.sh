#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_loop.py -m gpu -x -q 2>&1 | tail -5 > gpurun_out/pytest_loop.txt; cat gpurun_out/pytest_loop.txt
for wl in c3 c2; do
timeout 900 python bench.py --steps 24 --warmup 8 --no-cpu --workload $wl > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; tail -5 gpurun_out/bench_$wl.err
python -c "
import json; d=json.load(open('gpurun_out/bench_$wl.json')); print('$wl value %.1f Gcells/s  ms/step %.4f  frac %.3f whole %.3f' % (d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d['config']['whole_level_frac'])); print(' '.join('L%s:%.3f' % (k, v['hist_ms']) for k,v in d['config']['per_level'].items())); print(d['config']['phases_ms']); print(d['checksum']); print({k:v for k,v in d['e2e'].items() if k in ('value','ms_per_step','d2h_bytes_per_step')})"
done
