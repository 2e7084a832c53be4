#!/usr/bin/env python
"""BASELINE config 5 (the part that needs no Kubernetes): an end-to-end 100-tree bernoulli GBM on synthetic
50M x 100 through h2ok_hist.gbm.GBMTrainer on one GPU (or one rank per GPU under torchrun).  Prints one JSON line."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "h2o-kubernetes_b200"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from h2ok_hist import abi  # noqa: E402
from h2ok_hist.driver import CudaShard, EngineShard, LevelDriver, TreeLoop  # noqa: E402
from h2ok_hist.gbm import GBMTrainer  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=50_000_000)
ap.add_argument("--cols", type=int, default=100)
ap.add_argument("--trees", type=int, default=100)
ap.add_argument("--depth", type=int, default=5)  # H2O's GBM default
ap.add_argument("--bins", type=int, default=256)
ap.add_argument("--path", default="engine", choices=["stateless", "engine"])
a = ap.parse_args()
world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
row0 = a.rows * rank // world
rows = a.rows * (rank + 1) // world - row0
shard = (EngineShard if a.path == "engine" else CudaShard)(local)
if a.path == "engine":
    TreeLoop.join_ranks(shard)
shard.set_data_synth(abi.SHAPE_WIDE64, 20261106, row0, rows, a.cols)
# label = 1 if x0 + x2*x5 - x9 > tau; tau = the expression's median on a sample, so classes balance
import ctypes as C  # noqa: E402


def column(c):
    """Column c of this rank's rows, from the same counter-based generator that filled the shard."""
    t = torch.empty(shard.ld, dtype=torch.float32, device="cuda")
    shard.lib.call("h2ok_hist_synth_fill_cols_dev", shard.ctx, abi.SHAPE_WIDE64, 20261106, row0, rows, c, 1, shard.ld, 0,
                   C.c_void_p(t.data_ptr()))
    return t[:rows]


expr = column(0) + column(2) * column(5) - column(9)
tau = expr[: min(rows, 1_000_000)].median()
label = (expr > tau).float()
del expr
if a.path == "engine":  # the level loop inside the library
    drv = TreeLoop(shard, abi.Params.make(nbins=a.bins, max_depth=a.depth))
else:
    drv = LevelDriver(shard, abi.Params.make(nbins=a.bins, max_depth=a.depth))
    drv.prepare_data()
tr = GBMTrainer(drv, label, abi.DIST_BERNOULLI, learn_rate=0.1, keep_levels=False)
tr.init_f()
ll0 = tr.loss()
torch.cuda.synchronize()
t0 = time.perf_counter()
per_tree = []
for t in range(a.trees):
    t1 = time.perf_counter()
    tr.step()
    torch.cuda.synchronize()
    per_tree.append(time.perf_counter() - t1)
sec = time.perf_counter() - t0
ll = tr.loss()
if rank == 0:
    print(json.dumps({"workload": f"GBM bernoulli {a.trees} trees depth {a.depth}, {a.rows} x {a.cols}, {a.bins} bins",
                      "n_gpus": world, "path": a.path, "seconds": sec, "ms_per_tree": 1e3 * sec / a.trees,
                      "ms_per_tree_median": 1e3 * sorted(per_tree)[len(per_tree) // 2],
                      "cells_per_s_per_level": a.rows * a.cols * a.depth * a.trees / sec,
                      "logloss_start": ll0, "logloss_end": ll}))
if world > 1:
    dist.destroy_process_group()
