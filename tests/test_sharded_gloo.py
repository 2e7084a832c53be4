"""World-size-2 gloo tests of the N>1 path: rows sharded across ranks, the level's integer
histograms summed and the encoded observed ranges min/max-ed with torch.distributed, every rank deriving the same
splits.  The local compute is the oracle-backed CpuShard; what is under test is h2ok_hist.driver's host logic
(LevelDriver, reduce_level), which is the same code the GPUs run under NCCL."""
import os
import sys
import tempfile

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, init_file, split_rows, weight_mode, out_dir):
    for p in (os.path.join(ROOT, "h2o-kubernetes_b200"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import oracle_lib
    from cpu_shard import CpuShard
    from h2ok_hist import abi
    from h2ok_hist.driver import LevelDriver

    dist.init_process_group("gloo", init_method=f"file://{init_file}", rank=rank, world_size=world)
    oracle = oracle_lib.load()
    rows_total, cols, nb = 30000, 9, 64
    bounds = [0] + list(split_rows) + [rows_total]
    row0, row1 = bounds[rank], bounds[rank + 1]
    params = abi.Params.make(nbins=nb, max_depth=5)
    Xs, ys, ws = oracle.synth(abi.SHAPE_WIDE64, 55, row0, row1 - row0, cols, na_permille=10, weight_mode=weight_mode,
                              ld=max(4, (row1 - row0 + 3) // 4 * 4))
    drv = LevelDriver(CpuShard(oracle, Xs, row1 - row0, ys, ws), params)
    root = drv.prepare_data()
    levels = []

    def inspect(d, n, nch):
        sh = d.shard
        levels.append(dict(acc=sh.acc.numpy().copy(), splits=sh.splits.copy(), descs=sh.descs[:n].copy(),
                           min2=sh.min2.copy(), maxin=sh.maxin.copy(), ids=np.asarray(sh.child_ids).copy()))

    drv.train_tree(inspect=inspect)
    # unsharded reference computed locally by the same oracle
    X, y, w = oracle.synth(abi.SHAPE_WIDE64, 55, 0, rows_total, cols, na_permille=10, weight_mode=weight_mode)
    assert root.tobytes() == oracle.root_descs(X, rows_total, params).tobytes()
    q, _ = oracle.choose_quant(y, w, 25)
    assert (q.scale_wy, q.scale_wyy, q.scale_w) == (drv.shard.quant.scale_wy, drv.shard.quant.scale_wyy, drv.shard.quant.scale_w)
    wq, wyq, wyyq = oracle.quantize_rows(y, w, q)
    descs, nid, leaf, gid = root.reshape(1, cols), np.zeros(rows_total, np.int32), np.full(rows_total, -1, np.int32), 0
    for depth, lv in enumerate(levels):
        acc, min2, maxin = oracle.build_level_q(X, rows_total, nid, wq, wyq, wyyq, descs, nb)
        assert np.array_equal(lv["acc"], acc), f"rank {rank} level {depth}: reduced histogram != unsharded"
        assert np.array_equal(lv["min2"], min2) and np.array_equal(lv["maxin"], maxin)
        splits = oracle.find_splits(oracle.finalize(acc, q, nb), descs, params.min_rows, params.min_split_improvement)
        assert lv["splits"].tobytes() == splits.tobytes()
        if depth + 1 < params.max_depth:
            child, ids, nch = oracle.next_level(splits, descs, min2, maxin, params, depth)
        else:
            child, ids, nch = None, np.full(2 * descs.shape[0], -1, np.int32), 0
        assert np.array_equal(lv["ids"], ids)
        oracle.apply_splits(X, rows_total, splits, ids, gid, nid, leaf)
        gid += descs.shape[0]
        descs = child
    assert len(levels) == params.max_depth
    n_local = row1 - row0
    assert np.array_equal(drv.shard.leaf[:n_local], leaf[row0:row1])  # each rank routed its own rows identically
    dist.barrier()
    dist.destroy_process_group()
    open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")


@pytest.mark.parametrize("split_rows,weight_mode", [((15000,), 0), ((7,), 1), ((29996,), 0)])
def test_two_rank_sharded_tree_equals_unsharded(split_rows, weight_mode):
    with tempfile.TemporaryDirectory() as td:
        init_file = os.path.join(td, "init")
        mp.spawn(_worker, args=(2, init_file, split_rows, weight_mode, td), nprocs=2, join=True)
        assert os.path.exists(os.path.join(td, "ok0")) and os.path.exists(os.path.join(td, "ok1"))


def test_reduce_level_is_identity_without_process_group():
    from h2ok_hist.driver import reduce_level

    acc = torch.arange(12, dtype=torch.int64)
    mn = torch.tensor([5, -7], dtype=torch.int32)
    mx = torch.tensor([9, 3], dtype=torch.int32)
    reduce_level(acc, mn, mx)
    assert acc.tolist() == list(range(12)) and mn.tolist() == [5, -7] and mx.tolist() == [9, 3]
