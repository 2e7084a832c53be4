#!/bin/bash
# N = 2 and N = 4 lines on one 4-GPU box
mkdir -p gpurun_out
for n in 2 4; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $n --steps 32 --warmup 8 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; tail -3 gpurun_out/bench_n$n.err
python -c "
import json; d=json.load(open('gpurun_out/bench_n$n.json')); print('N=$n value %.1f Gcells/s  ms/step %.3f  frac %.3f whole %.3f' % (d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d['config']['whole_level_frac'])); print(d['config']['phases_ms']); print(d['checksum'], d['clocks']); print({k:v for k,v in d['e2e'].items() if k in ('value','ms_per_step','d2h_bytes_per_step')})"
done
