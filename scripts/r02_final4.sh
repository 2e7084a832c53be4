#!/bin/bash
# last check of the round: smoke(), the whole GPU suite, the default bench line and the reference arm
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 2400 python -m pytest tests -q -m gpu 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/r02_bench_c3_n1.json 2> gpurun_out/b1.err; tail -3 gpurun_out/b1.err
timeout 900 python bench.py --impl reference --steps 8 --warmup 3 > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/bref.err; tail -3 gpurun_out/bref.err
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_c3_n1.json')); print('c3 ms/step %.4f frac %.3f whole %.3f e2e %.3f launches %s' % (d['ms_per_step'], d['roofline']['frac'], d['config']['whole_level_frac'], d['e2e']['ms_per_step'], d['gpu_launches'])); print(d['config']['phases_ms']); print(d['checksum'], d['clocks']); print(d['cpu_baseline']['value'], d['cpu_baseline']['cores'], d['cpu_baseline']['checksum_slice_match'])"
cut -c1-300 gpurun_out/r02_bench_reference_arm.json
