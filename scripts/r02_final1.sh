#!/bin/bash
# final round-2 evidence on one GPU: the default bench line (with CPU baseline and e2e), the other workloads
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r02_bench_c3_n1.json 2> gpurun_out/b1.err; tail -3 gpurun_out/b1.err
for wl in c2 c4 c4-forest c3-u8; do
timeout 900 python bench.py --steps 30 --warmup 8 --no-cpu --workload $wl > gpurun_out/r02_bench_${wl}_n1.json 2> gpurun_out/b_$wl.err; tail -3 gpurun_out/b_$wl.err
done
timeout 900 python bench.py --impl reference --steps 8 --warmup 3 > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/bref.err; tail -3 gpurun_out/bref.err
for f in gpurun_out/r02_bench_c3_n1.json gpurun_out/r02_bench_c2_n1.json gpurun_out/r02_bench_c4_n1.json gpurun_out/r02_bench_c3-u8_n1.json; do
python -c "
import json,sys; d=json.load(open('$f')); print('$f'); print(' value %.1f Gcells/s  ms/step %.4f  frac %.3f whole %.3f' % (d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d['config']['whole_level_frac'])); print(' ', d['config']['phases_ms']); print(' ', d['checksum'], d['config'].get('fused_level_ms')); print(' ', {k:v for k,v in (d['e2e'] or {}).items() if k in ('value','ms_per_step','d2h_bytes_per_step','x_upload')}); print(' ', d.get('cpu_baseline')); print(' ', d['clocks'], d['gpu_launches'])"
done
cut -c1-400 gpurun_out/r02_bench_reference_arm.json
