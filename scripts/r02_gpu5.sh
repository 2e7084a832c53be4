#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --steps 16 --warmup 8 "$@" > gpurun_out/bench_tmp.json 2> gpurun_out/bench_tmp.err; tail -5 gpurun_out/bench_tmp.err
python -c "
import json; d=json.load(open('gpurun_out/bench_tmp.json')); print('value %.1f Gcells/s  ms/step %.3f  frac %.3f whole %.3f' % (d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d['config']['whole_level_frac'])); print(' '.join('L%s:%.2f' % (k, v['hist_ms']) for k,v in d['config']['per_level'].items())); print(d['config']['phases_ms']); print(d['checksum']); print(d['cpu_baseline']); print(d['e2e'])"
