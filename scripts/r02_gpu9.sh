#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_loop.py -m gpu -x -q -k "sparse" 2>&1 | tail -5
for wl in "$@"; do
timeout 1200 python bench.py --steps 30 --warmup 6 --no-cpu --no-e2e --workload $wl > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; tail -3 gpurun_out/bench_$wl.err
python -c "
import json; d=json.load(open('gpurun_out/bench_$wl.json')); print('$wl value %.1f Gcells/s  ms/step %.4f  frac %.3f whole %.3f' % (d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d['config']['whole_level_frac'])); print(' '.join('L%s:%.3f' % (k, v['hist_ms']) for k,v in d['config']['per_level'].items())); print(d['config']['phases_ms']); print(d['checksum']); print(d['config'].get('fused_level_ms'), d['config'].get('frac_of_dense_fused_floor'))"
done
