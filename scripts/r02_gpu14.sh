#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -12 > gpurun_out/pytest_gpu.txt; cat gpurun_out/pytest_gpu.txt
N=$(python -c "import torch; print(torch.cuda.device_count())")
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $N --steps 16 --warmup 8 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; tail -3 gpurun_out/bench_n$N.err
python -c "
import json; d=json.load(open('gpurun_out/bench_n$N.json')); print('N=$N value %.1f Gcells/s  ms/step %.3f  frac %.3f whole %.3f' % (d['value']/1e9, d['ms_per_step'], d['roofline']['frac'], d['config']['whole_level_frac'])); print(d['config']['phases_ms']); print(d['checksum']); print({k:v for k,v in d['e2e'].items() if k in ('value','ms_per_step','d2h_bytes_per_step')})"
