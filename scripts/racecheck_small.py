"""Small trees through every histogram / routing kernel, for `compute-sanitizer --tool racecheck`:
the row-wise engine (unweighted, weighted, with NaNs), the sparse column-list pass, and the stateless kernels."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "h2o-kubernetes_b200"))
import torch  # noqa: E402

from h2ok_hist import abi  # noqa: E402
from h2ok_hist.driver import CudaShard, EngineShard, LevelDriver, TreeLoop  # noqa: E402

params = abi.Params.make(nbins=64, max_depth=3, min_rows=2)
for name, sparse, shape, cols, na, wm in [("engine", False, abi.SHAPE_HIGGS28, 28, 0, 0), ("engine+na", False, abi.SHAPE_WIDE64, 70, 20, 0),
                                          ("engine weighted", False, abi.SHAPE_HIGGS28, 28, 10, 1), ("sparse", True, abi.SHAPE_SPARSE, 130, 0, 0)]:
    sh = EngineShard(0, sparse=sparse)
    sh.set_data_synth(shape, 7, 0, 6000, cols, na_permille=na, weight_mode=wm)
    loop = TreeLoop(sh, params, checksum=True)
    levels = loop.train_tree()
    torch.cuda.synchronize()
    print(name, "levels", levels, "checksum", ["%016x" % c for c in loop.checksum(0)], flush=True)
    sh.close()
sh = CudaShard(0)
sh.set_data_synth(abi.SHAPE_HIGGS28, 7, 0, 6000, 28, na_permille=10)
drv = LevelDriver(sh, params)
drv.prepare_data()
print("stateless levels", drv.train_tree(), flush=True)
torch.cuda.synchronize()
sh.close()
