#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_loop.py -x -q -m gpu -k "sparse" 2>&1 | tail -1
for pf in 2048 8192 32768; do
echo "== PF $pf"
H2OK_CSC_PF=$pf timeout 900 python bench.py --steps 12 --warmup 6 --no-cpu --workload c4 > gpurun_out/v.json 2> gpurun_out/v.err; tail -3 gpurun_out/v.err
python -c "
import json; d=json.load(open('gpurun_out/v.json')); print('ms/step %.4f' % d['ms_per_step']); print(' '.join('L%s(%d):%.2f' % (k, v['nodes'], v['hist_ms']) for k,v in d['config']['per_level'].items())); print(d['checksum'])"
done
