"""Worker for tests/test_multi_gpu.py (run under torch.distributed.run, one rank per GPU): a depth-6 tree on a
row-sharded synthetic matrix through LevelDriver + NCCL; rank 0 then builds the same tree alone on the whole matrix
and checks that every level's reduced integer histogram, ranges, splits and the leaf of every row are identical."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "h2o-kubernetes_b200"))

from h2ok_hist import abi  # noqa: E402
from h2ok_hist.driver import CudaShard, EngineShard, LevelDriver  # noqa: E402


def tree(shard, row0, rows, cols, params, weight_mode):
    shard.set_data_synth(abi.SHAPE_WIDE64, 4242, row0, rows, cols, na_permille=10, weight_mode=weight_mode)
    drv = LevelDriver(shard, params)
    drv.prepare_data()
    levels = []

    def inspect(d, n, nch):
        s = d.shard
        levels.append((n, nch, s.acc[:n].clone(), s.mn[:n].clone(), s.mx[:n].clone(), s.splits[:n].clone(),
                       s.child_ids[: 2 * n].clone()))

    assert drv.train_tree(inspect=inspect) == params.max_depth
    return levels, shard.leaf[:rows].clone()


def loop_tree(shard, row0, rows, cols, params, weight_mode, mode):
    """The same tree through the level loop inside the library (its own NCCL communicator, reduce mode `mode`)."""
    from h2ok_hist.driver import TreeLoop

    TreeLoop.join_ranks(shard)
    shard.set_data_synth(abi.SHAPE_WIDE64, 4242, row0, rows, cols, na_permille=10, weight_mode=weight_mode)
    loop = TreeLoop(shard, params, reduce_mode=mode, checksum=True)
    loop.begin_tree()
    levels = []
    while loop.n_nodes_left() > 0:
        loop.level()
        lv = loop.done_level()
        keep = dict(n=lv["n"], nch=lv["nch"], splits=lv["splits"].tobytes(), ids=lv["child_ids"].tobytes(),
                    min2=lv["min2"].tobytes(), maxin=lv["maxin"].tobytes(), pairs=(lv["pair0"], lv["pair1"]))
        # the accumulators this rank holds reduced: everything, or its share of the (node, column) pairs
        flat = lv["acc"].reshape(lv["n"] * cols, -1)
        keep["acc"] = flat[lv["pair0"]:lv["pair1"]].copy()
        levels.append(keep)
    return levels, shard.leaf[:rows].clone(), loop.checksum(row0)


def main_loop(mode_name, weight_mode):
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    mode = abi.REDUCE_ALL if mode_name == "loop-all" else abi.REDUCE_SCATTER
    total, cols = 400_003, 37
    params = abi.Params.make(nbins=64, max_depth=6, min_rows=4)
    row0 = total * rank // world
    rows = total * (rank + 1) // world - row0
    sh = EngineShard(local)
    levels, leaf, cks = loop_tree(sh, row0, rows, cols, params, weight_mode, mode)
    sh.close()
    parts = [torch.empty(total * (r + 1) // world - total * r // world, dtype=leaf.dtype, device=leaf.device)
             for r in range(world)]
    dist.all_gather(parts, leaf)
    gathered = [None] * world
    dist.all_gather_object(gathered, [(lv["pairs"], lv["acc"]) for lv in levels])
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        ref_levels, ref_leaf, ref_cks = loop_tree(EngineShard(local), 0, total, cols, params, weight_mode, mode)
        assert cks == ref_cks, (cks, ref_cks)
        assert len(levels) == len(ref_levels)
        for i, (a, b) in enumerate(zip(levels, ref_levels)):
            for k in ("n", "nch", "splits", "ids", "min2", "maxin"):
                assert a[k] == b[k], (i, k)
            full = b["acc"]
            for r in range(world):  # every rank's share of the reduced accumulators against the one-GPU histogram
                (p0, p1), acc = gathered[r][i]
                assert (acc == full[p0:p1]).all(), (i, r)
            if mode == abi.REDUCE_SCATTER:
                assert sum(g[i][0][1] - g[i][0][0] for g in gathered) == full.shape[0]
        assert torch.equal(torch.cat(parts), ref_leaf)
        print(f"MGPU_OK world={world} path={mode_name} weighted={weight_mode}")


def main():
    path, weight_mode = sys.argv[1], int(sys.argv[2])
    if path.startswith("loop"):
        return main_loop(path, weight_mode)
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    Shard = EngineShard if path == "engine" else CudaShard
    total, cols = 400_003, 37
    params = abi.Params.make(nbins=64, max_depth=6, min_rows=4)
    row0 = total * rank // world
    rows = total * (rank + 1) // world - row0
    levels, leaf = tree(Shard(local), row0, rows, cols, params, weight_mode)
    parts = [torch.empty(total * (r + 1) // world - total * r // world, dtype=leaf.dtype, device=leaf.device)
             for r in range(world)]
    dist.all_gather(parts, leaf)
    dist.barrier()
    dist.destroy_process_group()  # the single-GPU run below must not reduce
    if rank == 0:
        ref_levels, ref_leaf = tree(Shard(local), 0, total, cols, params, weight_mode)
        assert len(levels) == len(ref_levels)
        for a, b in zip(levels, ref_levels):
            assert a[0] == b[0] and a[1] == b[1]
            for ta, tb in zip(a[2:], b[2:]):
                assert torch.equal(ta, tb)
        assert torch.equal(torch.cat(parts), ref_leaf)
        print(f"MGPU_OK world={world} path={path} weighted={weight_mode}")


if __name__ == "__main__":
    main()
