"""CPU checks of bench.py's contract: the reference arm's JSON line (the CPU oracle port timed on the host cores, what
the driver pairs with the GPU arm's line) and the GPU arm's refusal to run without a device (no CPU fallback)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_bench(*args):
    env = dict(os.environ)
    env.pop("RANK", None)
    env.pop("WORLD_SIZE", None)
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], cwd=ROOT, env=env, capture_output=True,
                          text=True, timeout=600)


def test_reference_arm_line():
    p = run_bench("--impl", "reference", "--steps", "2", "--warmup", "1")
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [ln for ln in p.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1  # ONE JSON line on stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference"
    assert d["metric"] == "gbm_histogram_rows_x_cols_per_s_per_tree_level" and d["unit"] == "cells/s"
    assert d["higher_is_better"] is True and d["vs_baseline"] is None and d["data"] == "synthetic"
    assert d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] >= 3  # at least three warm-up steps whatever was asked
    assert d["value"] > 0 and d["ms_per_step"] > 0
    assert "100M rows x 64 cols" in d["config"]["workload"] and "model" not in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["sample"] and cb["value"] == d["value"]
    e = d["e2e"]
    assert e["value"] == d["value"] and e["unit"] == d["unit"]
    assert e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0


def test_gpu_arm_fails_loudly_without_a_device():
    import torch

    if torch.cuda.is_available():
        import pytest

        pytest.skip("a GPU is present")
    p = run_bench("--steps", "2", "--warmup", "1")
    assert p.returncode != 0
    assert not p.stdout.strip()  # no line at all rather than a line measured somewhere else
