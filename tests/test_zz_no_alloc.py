"""GPU test of h2ok_hist_next_level_noalloc_dev — the entry point for hosts that may not allocate device memory
while they run (a process whose GPU memory is entirely owned by a pool allocator, or a loop that cannot afford the
device-wide synchronisation of cudaMalloc/cudaFree).  It keeps next_level's scratch in a fixed device array inside
the library instead of a context buffer that grows on demand.

The file name sorts last on purpose: with `pytest -x` every other test has already run if this one fails.  It passes
on its own in a fresh process and after the rest of the suite, because everything it needs (two real tree levels) is
produced here from the oracle.  Both calls go through the C ABI; the results must be byte-identical to
h2ok_hist_next_level_dev (descriptors, child ids, child count are integer / exactly rounded fp64 work, so the
tolerance used everywhere else for them applies: none).  Two calls in a row, the second on a smaller level, show that
what the shared array keeps from the first call cannot leak into the second.
"""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")
from h2ok_hist import abi, capi  # noqa: E402

pytestmark = pytest.mark.gpu


def P(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def oracle_levels(oracle):
    """Two parent levels of a real tree (many nodes first, then fewer) with everything next_level consumes."""
    rows, cols = 40000, 12
    X, y, _ = oracle.synth(abi.SHAPE_WIDE64, 2024, 0, rows, cols, na_permille=10)
    params = abi.Params.make(nbins=64, max_depth=6, min_rows=5)
    levels, _, _ = oracle.train_tree(X, rows, y, None, params)
    deep, shallow = levels[4], levels[1]  # 16-ish nodes, then 2 nodes
    assert deep["descs"].shape[0] > shallow["descs"].shape[0] >= 2
    return params, cols, [deep, shallow]


def run(lib, ctx, name, params, cols, lv):
    n = lv["descs"].shape[0]
    d_splits = dev(lv["splits"].view(np.uint8))
    d_descs = dev(lv["descs"].view(np.uint8))
    d_min2, d_maxin = dev(lv["min2"]), dev(lv["maxin"])
    child = torch.full((2 * n, cols, abi.COL_DESC.itemsize), 0xAB, dtype=torch.uint8, device="cuda")
    ids = torch.full((2 * n,), -7, dtype=torch.int32, device="cuda")
    nch = torch.full((1,), -7, dtype=torch.int32, device="cuda")
    lib.call(name, ctx, P(d_splits), P(d_descs), P(d_min2), P(d_maxin), n, cols, C.byref(params), 0, P(child), P(ids), P(nch))
    torch.cuda.synchronize()
    k = int(nch.item())
    return child[:k].cpu().numpy().tobytes(), ids.cpu().numpy(), k


def test_no_alloc_next_level_matches_allocating_entry_point(oracle, lib):
    params, cols, levels = oracle_levels(oracle)
    ctx = C.c_void_p()
    lib.call("h2ok_hist_create", 0, C.byref(ctx))
    # the library's kernels go on the stream torch fills the output buffers on, so the fills cannot land late
    lib.call("h2ok_hist_set_stream", ctx, C.c_void_p(torch.cuda.current_stream().cuda_stream), 0)
    try:
        for lv in levels:  # two calls in a row through the shared scratch array
            want = run(lib, ctx, "h2ok_hist_next_level_dev", params, cols, lv)
            got = run(lib, ctx, "h2ok_hist_next_level_noalloc_dev", params, cols, lv)
            assert got[2] == want[2] and got[2] > 0
            assert np.array_equal(got[1], want[1])
            assert got[0] == want[0]
            # and both agree with the oracle's children
            child, ids, n = oracle.next_level(lv["splits"], lv["descs"], lv["min2"], lv["maxin"], params, 0)
            assert n == got[2] and np.array_equal(ids, got[1]) and child.tobytes() == got[0]
    finally:
        lib.call("h2ok_hist_destroy", ctx)


def test_no_alloc_next_level_rejects_levels_beyond_its_fixed_scratch(lib):
    ctx = C.c_void_p()
    lib.call("h2ok_hist_create", 0, C.byref(ctx))
    try:
        params = abi.Params.make(nbins=64)
        buf = torch.zeros(1 << 20, dtype=torch.uint8, device="cuda")
        for n_nodes, cols in [(129, 64), (1023, 9), (64, 129)]:  # 2*n_nodes*cols > 16384
            with pytest.raises(capi.H2okError) as e:
                lib.call("h2ok_hist_next_level_noalloc_dev", ctx, P(buf), P(buf), P(buf), P(buf), n_nodes, cols,
                         C.byref(params), 0, P(buf), P(buf), P(buf))
            assert e.value.code == abi.ERR_ARG and "H2OK_NOALLOC_MAX" in str(e.value)
    finally:
        lib.call("h2ok_hist_destroy", ctx)
