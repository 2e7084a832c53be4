#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_group.py -m gpu -x -q 2>&1 | tail -3
for n in 8; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $n --steps 32 --warmup 8 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; tail -3 gpurun_out/bench_n$n.err
python -c "
import json; d=json.load(open('gpurun_out/bench_n$n.json')); print('N=$n value %.1f Gcells/s  ms/step %.3f  frac %.3f whole %.3f' % (d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d['config']['whole_level_frac'])); print(d['config']['phases_ms']); print(d['config']['per_level']); print(d['checksum']); print({k:v for k,v in d['e2e'].items() if k in ('value','ms_per_step','d2h_bytes_per_step')})"
done
timeout 300 ./h2o-kubernetes_b200/tools/h2ok-bench --shape c3 --gpus 8 --trees 3 | tee gpurun_out/r02_h2ok_bench_c3_n8.json
