"""GPU tests of h2ok_hist_group_* — several devices behind ONE host thread, no torch anywhere on the data path — and
of the h2ok-bench binary that is built on it.  On a one-GPU box the group has one member; on a box with more GPUs the
same tree is built on all of them and must be the single-device tree, bit for bit (checksums of every level's reduced
accumulators, of the splits and of every row's leaf, plus the host buffers of every level)."""
import ctypes as C
import json
import os
import subprocess

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import oracle_lib  # noqa: E402
from h2ok_hist import abi, deploy  # noqa: E402

pytestmark = pytest.mark.gpu


def group_tree(lib, n_dev, shape, rows, cols, params, reduce_mode, want_out):
    devs = (C.c_int32 * n_dev)(*range(n_dev))
    g = C.c_void_p()
    lib.call("h2ok_hist_group_create", devs, n_dev, C.byref(g))
    levels = []
    try:
        assert lib.value("h2ok_hist_group_size", g) == n_dev
        lib.call("h2ok_hist_group_load_synth", g, shape, 20261106, rows, cols, 10)
        lib.call("h2ok_hist_group_loop_create", g, C.byref(params), None)
        lib.call("h2ok_hist_group_set_options", g, reduce_mode, 1, 0)
        lib.call("h2ok_hist_group_begin_tree_synth", g, 20261106, 0)
        nb1 = max(params.nbins, params.nbins_top_level) + 1
        n, nch = 1, C.c_int32(1)
        while nch.value > 0 and len(levels) < params.max_depth:
            out = None
            if want_out:
                bufs = dict(hist=np.full((n, cols, nb1, 3), np.nan), min2=np.zeros((n, cols), np.float32),
                            maxin=np.zeros((n, cols), np.float32), splits=np.zeros(n, abi.SPLIT),
                            child_descs=np.zeros((2 * n, cols), abi.COL_DESC), child_ids=np.zeros(2 * n, np.int32))
                out = abi.LevelOut(*[C.c_void_p(bufs[k].ctypes.data) for k in ("hist", "min2", "maxin", "splits", "child_descs", "child_ids")])
            done = C.c_int32()
            lib.call("h2ok_hist_group_level", g, None if out is None else C.byref(out), C.byref(done), C.byref(nch))
            assert done.value == n
            if want_out:
                bufs["child_descs"] = bufs["child_descs"][: nch.value]
                levels.append(bufs)
            else:
                levels.append(None)
            n = nch.value
        sums = (C.c_uint64 * 3)()
        lib.call("h2ok_hist_group_checksum", g, sums)
        ms = C.c_float()
        lib.call("h2ok_hist_group_hist_ms", g, C.byref(ms))
        assert ms.value > 0
        return [int(x) for x in sums], levels
    finally:
        lib.call("h2ok_hist_group_destroy", g)


def test_group_of_one_matches_the_oracle(lib, oracle):
    rows, cols = 60007, 28
    params = abi.Params.make(nbins=64, max_depth=5)
    sums, levels = group_tree(lib, 1, abi.SHAPE_HIGGS28, rows, cols, params, abi.REDUCE_SCATTER, True)
    X, y, _ = oracle.synth(abi.SHAPE_HIGGS28, 20261106, 0, rows, cols, na_permille=10)
    ref, _, leaf, _ = oracle_lib.train_tree_q(oracle, X, rows, y, None, params)
    assert sums == oracle_lib.tree_checksum(ref, leaf)
    assert len(levels) == len(ref)
    for got, lv in zip(levels, ref):
        assert np.array_equal(got["hist"], lv["hist"])
        assert np.array_equal(got["min2"], lv["min2"]) and np.array_equal(got["maxin"], lv["maxin"])
        assert got["splits"].tobytes() == lv["splits"].tobytes()
        assert np.array_equal(got["child_ids"], lv["child_ids"])
        assert got["child_descs"].tobytes() == lv["child_descs"].tobytes()


@pytest.mark.parametrize("mode", [abi.REDUCE_SCATTER, abi.REDUCE_ALL])
def test_group_of_all_gpus_is_bit_identical_to_one(lib, mode):
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs two GPUs")
    rows, cols = 500_009, 37
    params = abi.Params.make(nbins=64, max_depth=6, min_rows=4)
    one, lv1 = group_tree(lib, 1, abi.SHAPE_WIDE64, rows, cols, params, mode, True)
    many, lvn = group_tree(lib, n, abi.SHAPE_WIDE64, rows, cols, params, mode, True)
    assert one == many
    assert len(lv1) == len(lvn)
    for a, b in zip(lv1, lvn):  # one set of host buffers, filled by all members: complete and identical
        for k in a:
            assert a[k].tobytes() == b[k].tobytes(), k



def group_tree_host(lib, n_dev, X, rows, y, params, is_int):
    """One tree over n_dev devices on the caller's columns (h2ok_hist_group_set_data_host / begin_tree_host)."""
    cols = X.shape[0]
    devs = (C.c_int32 * n_dev)(*range(n_dev))
    g = C.c_void_p()
    lib.call("h2ok_hist_group_create", devs, n_dev, C.byref(g))
    try:
        ptrs = (C.c_void_p * cols)(*[X[c].ctypes.data for c in range(cols)])
        lib.call("h2ok_hist_group_set_data_host", g, ptrs, rows, cols)
        ii = np.asarray(is_int, np.int32)
        lib.call("h2ok_hist_group_loop_create", g, C.byref(params), abi.ptr(ii))
        lib.call("h2ok_hist_group_set_options", g, abi.REDUCE_SCATTER, 1, 0)  # (categorical columns: ALL is used anyway)
        lib.call("h2ok_hist_group_begin_tree_host", g, abi.ptr(y), None)
        splits = []
        n, nch = 1, C.c_int32(1)
        while nch.value > 0 and len(splits) < params.max_depth:
            sp = np.zeros(n, abi.SPLIT)
            out = abi.LevelOut(None, None, None, C.c_void_p(sp.ctypes.data), None, None)
            done = C.c_int32()
            lib.call("h2ok_hist_group_level", g, C.byref(out), C.byref(done), C.byref(nch))
            splits.append(sp)
            n = nch.value
        sums = (C.c_uint64 * 3)()
        lib.call("h2ok_hist_group_checksum", g, sums)
        return [int(x) for x in sums], splits
    finally:
        lib.call("h2ok_hist_group_destroy", g)


def test_group_categorical_columns_match_the_oracle_on_every_device_count(lib, oracle):
    """Bit-set splits through the group API: one device against the oracle, and all the box's devices against one."""
    from test_loop import categorical_case

    rows, cols = 80021, 9
    X, y, is_int = categorical_case(oracle, rows, cols, 4242, 10)
    X = np.ascontiguousarray(X[:, :rows])
    y = np.ascontiguousarray(y[:rows])
    params = abi.Params.make(nbins=512, max_depth=6, min_rows=5)
    ref, _, leaf, _ = oracle_lib.train_tree_q(oracle, X, rows, y, None, params, is_int)
    want = oracle_lib.tree_checksum(ref, leaf)
    assert any((lv["splits"]["equal"] == 2).any() for lv in ref)
    for n in sorted({1, torch.cuda.device_count()}):
        sums, splits = group_tree_host(lib, n, X, rows, y, params, is_int)
        assert sums == want, n
        assert len(splits) == len(ref)
        for got, lv in zip(splits, ref):
            assert got.tobytes() == lv["splits"].tobytes()

def test_h2ok_bench_binary_runs_and_agrees_with_the_library(lib):
    rows, cols = 200_000, 28
    out = subprocess.run([deploy.BENCH_BINARY, "--shape", "c2", "--rows", str(rows), "--bins", "64", "--depth", "5", "--trees", "2"],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    line = json.loads(out.stdout)  # stdout is a pipe here: exactly one JSON line
    assert line["n_gpus"] == 1 and line["levels"] == 10 and line["value"] > 0
    params = abi.Params.make(nbins=64, max_depth=5)
    devs = (C.c_int32 * 1)(0)
    g = C.c_void_p()
    lib.call("h2ok_hist_group_create", devs, 1, C.byref(g))
    try:
        lib.call("h2ok_hist_group_load_synth", g, abi.SHAPE_HIGGS28, 20261106, rows, cols, 0)
        lib.call("h2ok_hist_group_loop_create", g, C.byref(params), None)
        lib.call("h2ok_hist_group_set_options", g, abi.REDUCE_SCATTER, 1, 0)
        lib.call("h2ok_hist_group_begin_tree_synth", g, 20261106, 0)
        nch = C.c_int32(1)
        while nch.value:
            lib.call("h2ok_hist_group_level", g, None, None, C.byref(nch))
        sums = (C.c_uint64 * 3)()
        lib.call("h2ok_hist_group_checksum", g, sums)
    finally:
        lib.call("h2ok_hist_group_destroy", g)
    assert line["checksum"] == ["%016x" % int(s) for s in sums]
