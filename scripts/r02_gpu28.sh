#!/bin/bash
mkdir -p gpurun_out
for v in 0 1 2; do
echo "== eng variant $v"
H2OK_ENG_VARIANT=$v timeout 900 python -m pytest tests/test_loop.py -x -q -m gpu -k "test_loop_levels_match_oracle or integer_columns" 2>&1 | tail -1
H2OK_ENG_VARIANT=$v timeout 900 python bench.py --steps 16 --warmup 8 --no-cpu --no-e2e > gpurun_out/e$v.json 2> gpurun_out/e$v.err; tail -3 gpurun_out/e$v.err
python -c "
import json; d=json.load(open('gpurun_out/e$v.json')); print('ms/step %.4f frac %.3f clk %s' % (d['ms_per_step'], d['roofline']['frac'], d['clocks']['sm_mhz'])); print(' '.join('L%s:%.2f' % (k, v['hist_ms']) for k,v in d['config']['per_level'].items())); print(d['checksum'])"
done
