#!/bin/bash
# after the scheduling barriers in the dense histogram pass: the whole GPU suite, then the lines of the dense workloads
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -q -m gpu 2>&1 | tail -4
timeout 900 python bench.py > gpurun_out/r02_bench_c3_n1.json 2> gpurun_out/b1.err; tail -3 gpurun_out/b1.err
for wl in c2 c3-u8; do
timeout 900 python bench.py --steps 30 --warmup 8 --no-cpu --workload $wl > gpurun_out/r02_bench_${wl}_n1.json 2> gpurun_out/b_$wl.err; tail -3 gpurun_out/b_$wl.err
done
for f in gpurun_out/r02_bench_c3_n1.json gpurun_out/r02_bench_c2_n1.json gpurun_out/r02_bench_c3-u8_n1.json; do
python -c "
import json,sys; d=json.load(open('$f')); print('$f'); print(' value %.1f Gcells/s  ms/step %.4f  frac %.3f whole %.3f' % (d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d['config']['whole_level_frac'])); print(' ', d['config']['phases_ms']); print(' ', ' '.join('L%s:%.2f' % (k, v['hist_ms']) for k,v in d['config']['per_level'].items())); print(' ', d['checksum']); print(' ', {k:v for k,v in (d['e2e'] or {}).items() if k in ('value','ms_per_step')}); print(' ', d['clocks'], d['gpu_launches'])"
done
