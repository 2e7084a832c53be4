#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scripts/gbm_c4.py --iters 2 --storage sparse > gpurun_out/r02_gbm_c4_sparse.json 2> gpurun_out/gbm_c4_sparse.err; tail -2 gpurun_out/gbm_c4_sparse.err; cat gpurun_out/r02_gbm_c4_sparse.json
timeout 900 python scripts/gbm_c4.py --iters 2 --storage dense > gpurun_out/r02_gbm_c4_dense.json 2> gpurun_out/gbm_c4_dense.err; tail -2 gpurun_out/gbm_c4_dense.err; cat gpurun_out/r02_gbm_c4_dense.json
timeout 900 python scripts/gbm_c5.py > gpurun_out/r02_gbm_c5.json 2> gpurun_out/gbm_c5.err; tail -2 gpurun_out/gbm_c5.err; cat gpurun_out/r02_gbm_c5.json
timeout 900 python bench.py --steps 30 --warmup 6 --no-cpu --no-e2e --workload c4-dense > gpurun_out/r02_bench_c4_dense_n1.json 2> gpurun_out/c4d.err; tail -2 gpurun_out/c4d.err; cut -c1-300 gpurun_out/r02_bench_c4_dense_n1.json
timeout 600 ./h2o-kubernetes_b200/tools/h2ok-bench --shape c3 --trees 2 | tee gpurun_out/r02_h2ok_bench_c3.json
