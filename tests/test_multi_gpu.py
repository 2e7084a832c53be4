"""N GPUs give the same bits as one: needs at least two visible GPUs, skipped otherwise."""
import os
import subprocess
import sys

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


PORTS = {"engine": 0, "stateless": 7, "loop-scatter": 14, "loop-all": 21}


@pytest.mark.parametrize("path,weight_mode", [("engine", 0), ("engine", 1), ("stateless", 0), ("loop-scatter", 0),
                                              ("loop-scatter", 1), ("loop-all", 0)])
def test_sharded_tree_is_bit_identical_to_one_gpu(path, weight_mode):
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs two GPUs")
    world = 4 if n >= 4 else 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(29600 + weight_mode + PORTS[path]),
           os.path.join(HERE, "mgpu_worker.py"), path, str(weight_mode)]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-3000:]
    assert f"MGPU_OK world={world} path={path}" in out.stdout + out.stderr
