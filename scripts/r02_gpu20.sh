#!/bin/bash
mkdir -p gpurun_out
for ps in 20 40 80; do
echo "== per_sm $ps"
H2OK_CSC_PER_SM=$ps H2OK_CSC_VARIANT=1 timeout 900 python bench.py --steps 12 --warmup 6 --no-cpu --workload c4 > gpurun_out/p$ps.json 2> gpurun_out/p$ps.err; tail -3 gpurun_out/p$ps.err
python -c "
import json; d=json.load(open('gpurun_out/p$ps.json')); print('ms/step %.4f' % d['ms_per_step']); print(' '.join('L%s(%d):%.2f' % (k, v['nodes'], v['hist_ms']) for k,v in d['config']['per_level'].items())); print(d['checksum'])"
done
