#!/usr/bin/env python
"""BASELINE config 4 end to end: K-class (multinomial) GBM on synthetic 10M x 2000 sparse-ish columns, 64 bins —
K one-class trees per iteration on softmax residuals through h2ok_hist.gbm.GBMTrainer on the level loop inside the
library (one GPU, or one rank per GPU under torchrun).  --storage sparse (default): column lists of nonzeros
(H2OK_TREE_SPARSE); dense: the row-wise engine on fp32, as round 1 ran it.  Prints one JSON line."""
import argparse
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "h2o-kubernetes_b200"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from h2ok_hist import abi  # noqa: E402
from h2ok_hist.driver import EngineShard, TreeLoop  # noqa: E402
from h2ok_hist.gbm import GBMTrainer  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=10_000_000)
ap.add_argument("--cols", type=int, default=2000)
ap.add_argument("--classes", type=int, default=10)
ap.add_argument("--iters", type=int, default=3)
ap.add_argument("--depth", type=int, default=6)
ap.add_argument("--bins", type=int, default=64)
ap.add_argument("--storage", default="sparse", choices=["sparse", "dense"])
ap.add_argument("--forest", action="store_true", help="build the K class trees of an iteration in lockstep (sparse storage)")
a = ap.parse_args()
world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
row0 = a.rows * rank // world
rows = a.rows * (rank + 1) // world - row0
shard = EngineShard(local, sparse=a.storage == "sparse")
TreeLoop.join_ranks(shard)
shard.set_data_synth(abi.SHAPE_SPARSE, 20261106, row0, rows, a.cols)


def column(c):
    t = torch.empty(shard.ld, dtype=torch.float32, device="cuda")
    shard.lib.call("h2ok_hist_synth_fill_cols_dev", shard.ctx, abi.SHAPE_SPARSE, 20261106, row0, rows, c, 1, shard.ld, 0,
                   C.c_void_p(t.data_ptr()))
    return t[:rows]


# class = which of the first K columns is largest (ties -> lowest index): a target the sparse columns can explain
score = torch.stack([column(c) for c in range(a.classes)])
label = score.argmax(dim=0).float()
del score
drv = TreeLoop(shard, abi.Params.make(nbins=a.bins, max_depth=a.depth), forest=a.classes if a.forest else 1)
tr = GBMTrainer(drv, label, abi.DIST_MULTINOMIAL, learn_rate=0.3, n_classes=a.classes, keep_levels=False)
tr.init_f()
ll0 = tr.loss()
torch.cuda.synchronize()
t0 = time.perf_counter()
per_iter = []
for _ in range(a.iters):
    t1 = time.perf_counter()
    tr.step()
    torch.cuda.synchronize()
    per_iter.append(time.perf_counter() - t1)
sec = time.perf_counter() - t0
ll = tr.loss()
levels = len(tr.trees) * a.depth  # upper bound: a class tree may stop early
if rank == 0:
    print(json.dumps({"workload": f"GBM multinomial {a.classes} classes, {a.iters} iterations ({len(tr.trees)} trees) depth {a.depth}, "
                                  f"{a.rows} x {a.cols}, {a.bins} bins",
                      "n_gpus": world, "path": "in-library loop, %s storage%s" % (a.storage, ", class trees as one forest" if a.forest else ""), "seconds": sec, "s_per_iteration": sec / a.iters,
                      "ms_per_tree": 1e3 * sec / len(tr.trees), "ms_per_level": 1e3 * sec / levels,
                      "cells_per_s_per_level": a.rows * a.cols * levels / sec,
                      "logloss_start": ll0, "logloss_end": ll}))
if world > 1:
    dist.destroy_process_group()
