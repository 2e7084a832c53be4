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


def main():
    path, weight_mode = sys.argv[1], int(sys.argv[2])
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
