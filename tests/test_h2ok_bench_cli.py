"""`h2ok bench` as delivered: the C++ binary tools/h2ok-bench.  No GPU needed here — help
text, flag validation and the hand-over from the Python `h2ok bench` parser are host logic.  The patterns are the ones
the reference's own integration tests use for its sub-commands (tests/integration_tests.rs:4, 32, 45-46) and the
messages of its validators (src/cli/mod.rs:245-252)."""
import os
import re
import subprocess

import pytest

from h2ok_hist import deploy

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = deploy.BENCH_BINARY


@pytest.fixture(scope="module", autouse=True)
def built():
    if not os.path.exists(BIN):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "h2o-kubernetes_b200", "csrc")], stdout=subprocess.DEVNULL)
    assert os.path.exists(BIN)


def run(*args):
    return subprocess.run([BIN, *args], capture_output=True, text=True, timeout=60)


def test_help_reads_like_a_subcommand_of_the_reference():
    out = run("-h")
    assert out.returncode == 0
    # the reference matches sub-command help with r"h2ok-deploy.*" and r"h2ok-undeploy.*\nUndeploys an existing ..."
    assert re.match(r"h2ok-bench.*\nBenchmarks the GBM histogram path.*", out.stdout)
    assert re.search(r"USAGE:\n    h2ok bench \[OPTIONS\]", out.stdout)
    for flag in ("--shape", "--rows", "--cols", "--bins", "--depth", "--gpus", "--trees", "--seed"):
        assert flag in out.stdout
    assert run("--help").stdout == out.stdout
    assert re.match(r"h2ok-bench \d+\.\d+\.\d+", run("-V").stdout)


@pytest.mark.parametrize("args,message", [
    (["--gpus", "0"], "Error: The number provided must be greater than zero."),  # cli/mod.rs:248
    (["--trees", "-3"], "Error: The number provided must be greater than zero."),
    (["--rows", "-1"], "must not be negative"),
    (["--shape", "c9"], "must be one of c2, c3, c4"),
    (["--bins", "5000"], "at most 1024 bins"),
])
def test_validators_reject_bad_values_like_clap(args, message):
    out = run(*args)
    assert out.returncode == 1 and out.stdout == ""
    assert out.stderr.startswith(f"error: Invalid value for '{args[0]} <{args[0][2:]}>'") and message in out.stderr


def test_unknown_flags_and_missing_values():
    out = run("--nope")
    assert out.returncode == 1 and "wasn't expected" in out.stderr and "USAGE" in out.stderr
    out = run("--rows")
    assert out.returncode == 1 and "requires a value" in out.stderr


def test_python_bench_subcommand_hands_over_to_the_binary():
    args = deploy.build_parser().parse_args(["bench", "--shape", "c2", "-g", "2", "--rows", "5000", "-t", "1"])
    argv = deploy.bench_argv(args)
    assert argv[0] == BIN and argv[1:] == ["--shape", "c2", "--rows", "5000", "--cols", "0", "--bins", "0", "--depth", "0",
                                            "--gpus", "2", "--trees", "1", "--seed", "20261106"]
    # without a GPU the binary fails loudly with the library's message, not with a CPU fallback
    import torch

    if not torch.cuda.is_available():
        out = subprocess.run(argv, capture_output=True, text=True, timeout=120)
        assert out.returncode == 1 and "h2ok_hist_group_create" in out.stderr


def test_rust_patch_is_present_and_marked_unverified():
    base = os.path.join(ROOT, "reference_patch")
    for rel in ("README.md", "h2ok_gpus.patch", "src/cli/bench.rs", "src/hist_ffi.rs", "src/k8s/templates_gpu.rs"):
        txt = open(os.path.join(base, rel)).read()
        assert "UNVERIFIED" in txt, rel
    patch = open(os.path.join(base, "h2ok_gpus.patch")).read()
    assert "#[serde(default)]" in patch and "Command::Bench" in patch and "<gpu-resources>" in patch
    # the flag table of the Rust source and of the binary agree
    rs = open(os.path.join(base, "src/cli/bench.rs")).read()
    for flag in ("shape", "rows", "cols", "bins", "depth", "gpus", "trees", "seed"):
        assert f'.long("{flag}")' in rs
