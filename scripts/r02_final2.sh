#!/bin/bash
# two GPUs: the whole GPU suite (multi-GPU tests included), then the config-4 lines and the default line on GPU 0
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -q -m gpu 2>&1 | tail -6
export CUDA_VISIBLE_DEVICES=0
for wl in c4 c4-forest; do
timeout 900 python bench.py --steps 18 --warmup 6 --no-cpu --workload $wl > gpurun_out/r02_bench_${wl}_n1.json 2> gpurun_out/b_$wl.err; tail -3 gpurun_out/b_$wl.err
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_${wl}_n1.json')); print('$wl value %.1f Gcells/s  ms/step %.4f  frac %.3f whole %.3f' % (d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d['config']['whole_level_frac'])); print(' '.join('L%s(%d):%.2f' % (k, v['nodes'], v['hist_ms']) for k,v in d['config']['per_level'].items())); print(d['config']['phases_ms']); print(d['checksum'], d['config'].get('fused_level_ms'), d['config'].get('frac_of_dense_fused_floor'), d.get('e2e', {}) and d['e2e'].get('ms_per_step'))"
done
timeout 900 python scripts/gbm_c4.py --iters 2 --storage sparse --forest > gpurun_out/r02_gbm_c4_forest.json 2> gpurun_out/gbm_c4_forest.err; tail -2 gpurun_out/gbm_c4_forest.err; cat gpurun_out/r02_gbm_c4_forest.json
timeout 900 python scripts/gbm_c4.py --iters 2 --storage sparse > gpurun_out/r02_gbm_c4_seq.json 2> gpurun_out/gbm_c4_seq.err; tail -2 gpurun_out/gbm_c4_seq.err; cat gpurun_out/r02_gbm_c4_seq.json
timeout 900 python bench.py > gpurun_out/r02_bench_c3_n1.json 2> gpurun_out/b_c3.err; tail -3 gpurun_out/b_c3.err
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_c3_n1.json')); print('c3 ms/step %.4f frac %.3f whole %.3f e2e %.3f' % (d['ms_per_step'], d['roofline']['frac'], d['config']['whole_level_frac'], d['e2e']['ms_per_step'])); print(d['config']['phases_ms']); print(d['checksum'], d['cpu_baseline'])"
