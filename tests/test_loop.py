"""GPU parity tests of the level loop that runs inside the library (h2ok_hist_loop_*): every level's descriptors,
integer accumulators, observed ranges, splits, child ids / descriptors, the row -> node map and the leaf of every row
must equal the oracle's fixed-point level loop bit for bit — on small ragged cases and at the shapes BASELINE.json
names (config 2 at full size, a 10M-row slice of config 3, config 4's sparse 2000-column shape)."""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import oracle_lib  # noqa: E402
from h2ok_hist import abi  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eshard():
    from h2ok_hist.driver import EngineShard

    s = EngineShard(0)
    yield s
    s.close()


@pytest.fixture(scope="module")
def sshard():
    from h2ok_hist.driver import EngineShard

    s = EngineShard(0, sparse=True)
    yield s
    s.close()


def check_tree(eshard, oracle, X, rows, y, w, params, is_int=None, host_out=False):
    from h2ok_hist.driver import TreeLoop

    cols = X.shape[0]
    eshard.set_data_tensors(torch.from_numpy(X[:, :rows].copy()), torch.from_numpy(y[:rows].copy()),
                            None if w is None else torch.from_numpy(w[:rows].copy()))
    loop = TreeLoop(eshard, params, is_int=is_int, checksum=True)
    assert loop.root_descs().tobytes() == oracle.root_descs(X, rows, params, is_int).tobytes()
    ref_levels, ref_nid, ref_leaf, q = oracle_lib.train_tree_q(oracle, X, rows, y, w, params, is_int)
    nb1 = loop.max_bins + 1
    loop.begin_tree()
    for lv in ref_levels:
        n = lv["descs"].shape[0]
        assert loop.n_nodes == n
        assert np.array_equal(eshard.nid[:rows].cpu().numpy(), lv["nid"])
        out = None
        if host_out:
            nch_max = 2 * n
            bufs = dict(hist=np.full((n, cols, nb1, 3), -1.0), min2=np.zeros((n, cols), np.float32),
                        maxin=np.zeros((n, cols), np.float32), splits=np.zeros(n, abi.SPLIT),
                        child_descs=np.zeros((nch_max, cols), abi.COL_DESC), child_ids=np.zeros(2 * n, np.int32))
            out = abi.LevelOut(*[C.c_void_p(bufs[k].ctypes.data) for k in ("hist", "min2", "maxin", "splits", "child_descs", "child_ids")])
        done, nch = loop.level(out)
        got = loop.done_level()
        assert done == n and got["n"] == n and got["depth"] == lv["depth"] and got["gid_base"] == lv["gid_base"]
        assert (got["pair0"], got["pair1"]) == (0, n * cols)
        assert got["descs"].tobytes() == lv["descs"].tobytes()
        assert np.array_equal(got["acc"], lv["acc"])
        assert np.array_equal(got["min2"], lv["min2"]) and np.array_equal(got["maxin"], lv["maxin"])
        assert got["splits"].tobytes() == lv["splits"].tobytes()
        assert np.array_equal(got["child_ids"], lv["child_ids"])
        assert nch == lv["child_descs"].shape[0]
        if nch:
            assert got["child_descs"].tobytes() == lv["child_descs"].tobytes()
        if host_out:
            assert np.array_equal(bufs["hist"], lv["hist"])
            assert np.array_equal(bufs["min2"], lv["min2"]) and np.array_equal(bufs["maxin"], lv["maxin"])
            assert bufs["splits"].tobytes() == lv["splits"].tobytes()
            assert np.array_equal(bufs["child_ids"], lv["child_ids"])
            assert bufs["child_descs"][:nch].tobytes() == lv["child_descs"].tobytes()
    assert loop.n_nodes_left() == 0
    assert np.array_equal(eshard.nid[:rows].cpu().numpy(), ref_nid)
    assert np.array_equal(eshard.leaf[:rows].cpu().numpy(), ref_leaf)
    assert loop.checksum(0) == oracle_lib.tree_checksum(ref_levels, ref_leaf)
    return ref_levels


@pytest.mark.parametrize("rows,cols,nbins,depth,weight_mode,na,host_out",
                         [(1, 3, 16, 3, 0, 0, False), (130, 4, 32, 5, 0, 20, True), (4097, 5, 32, 6, 1, 10, False),
                          (60001, 9, 64, 7, 0, 10, True), (50021, 40, 20, 5, 1, 10, False), (20011, 130, 32, 4, 0, 15, False),
                          (30011, 70, 256, 5, 0, 0, True)])
def test_loop_levels_match_oracle(eshard, oracle, rows, cols, nbins, depth, weight_mode, na, host_out):
    X, y, w = oracle.synth(abi.SHAPE_WIDE64, 500 + rows, 0, rows, cols, na_permille=na, weight_mode=weight_mode)
    top = 1024 if nbins == 20 else None  # H2O's defaults: nbins 20, nbins_top_level 1024 halved per level
    params = abi.Params.make(nbins=nbins, nbins_top_level=top, max_depth=depth, min_rows=1 if rows < 1000 else 5)
    check_tree(eshard, oracle, X, rows, y, w, params, host_out=host_out)


def test_loop_integer_columns_and_early_leaves(eshard, oracle):
    rows, cols = 50001, 10
    X, y, w = oracle.synth(abi.SHAPE_WIDE64, 99, 0, rows, cols, na_permille=30, weight_mode=1)
    X[4, :rows] = np.floor(X[4, :rows] * 3)
    X[7, :rows] = np.floor(X[7, :rows] * 1000)
    is_int = [0, 0, 0, 0, 1, 0, 0, 1, 0, 0]
    levels = check_tree(eshard, oracle, X, rows, y, w, abi.Params.make(nbins=64, max_depth=9, min_rows=50), is_int)
    assert len(levels) >= 5


def test_loop_config2_full_size_matches_oracle(eshard, oracle):
    """BASELINE config 2 as specified: 1M x 28 HIGGS-shaped, 256 bins, depth 6 — every level against the oracle."""
    rows, cols = 1_000_000, 28
    X, y, _ = oracle.synth(abi.SHAPE_HIGGS28, 20261106, 0, rows, cols)
    levels = check_tree(eshard, oracle, X, rows, y, None, abi.Params.make(nbins=256, max_depth=6))
    assert [lv["descs"].shape[0] for lv in levels] == [1, 2, 4, 8, 16, 32]


def test_loop_config3_slice_matches_oracle(eshard, oracle):
    """A 10M-row slice of BASELINE config 3 (64 columns, 256 bins, depth 8: 128 nodes at the last level)."""
    rows, cols = 10_000_000, 64
    X, y, _ = oracle.synth(abi.SHAPE_WIDE64, 20261106, 0, rows, cols)
    levels = check_tree(eshard, oracle, X, rows, y, None, abi.Params.make(nbins=256, max_depth=8))
    assert levels[-1]["descs"].shape[0] == 128


def test_loop_config4_shape_matches_oracle(eshard, oracle):
    """BASELINE config 4's shape: 2000 sparse-ish columns (90 % exact zeros), 64 bins — 63 column blocks."""
    rows, cols = 120_000, 2000
    X, y, _ = oracle.synth(abi.SHAPE_SPARSE, 20261106, 0, rows, cols)
    levels = check_tree(eshard, oracle, X, rows, y, None, abi.Params.make(nbins=64, max_depth=5))
    assert len(levels) == 5


@pytest.mark.parametrize("rows,cols,nbins,depth,na", [(5, 3, 16, 3, 0), (4099, 70, 32, 5, 20), (60001, 130, 64, 6, 5),
                                                      (30011, 64, 256, 5, 0), (20000, 200, 128, 4, 0), (50001, 9, 256, 8, 10)])
def test_sparse_storage_matches_oracle(sshard, oracle, rows, cols, nbins, depth, na):
    """H2OK_TREE_SPARSE: nonzero-only storage and histogram pass, zeros accounted for afterwards — same bits as the
    oracle at every level, on 90 %-zero data with NaNs, dense stretches, and negative zeros."""
    X, y, _ = oracle.synth(abi.SHAPE_SPARSE, 700 + rows, 0, rows, cols, na_permille=na)
    X[min(2, cols - 1), : rows // 2] = 1.5 + np.arange(rows // 2, dtype=np.float32) % 7  # a dense stretch of one column
    if cols > 64:
        X[64:min(cols, 96), : rows // 3] += np.float32(0.25)  # a block of columns without zeros for a third of the rows
    X[0, :rows] = np.where(np.arange(rows) % 5 == 0, -0.0, X[0, :rows])  # negative zeros are zeros
    params = abi.Params.make(nbins=nbins, max_depth=depth, min_rows=1 if rows < 1000 else 5)
    check_tree(sshard, oracle, X, rows, y, None, params, host_out=(rows % 2 == 1))
    assert sshard.stats().range_errors == 0


def test_sparse_storage_all_zero_and_all_dense_columns(sshard, oracle):
    rows, cols = 9001, 66
    X, y, _ = oracle.synth(abi.SHAPE_WIDE64, 3, 0, rows, cols)  # no zeros at all except the columns set below
    X[5, :rows] = 0.0
    X[65, :rows] = 0.0
    X[7, :rows] = np.where(np.arange(rows) % 2 == 0, 0.0, -1.0)
    levels = check_tree(sshard, oracle, X, rows, y, None, abi.Params.make(nbins=32, max_depth=4))
    assert len(levels) == 4


def test_sparse_config4_shape_matches_oracle(sshard, oracle):
    """BASELINE config 4's shape on the storage made for it: 2000 columns, 90 % exact zeros, 64 bins."""
    rows, cols = 120_000, 2000
    X, y, _ = oracle.synth(abi.SHAPE_SPARSE, 20261106, 0, rows, cols)
    levels = check_tree(sshard, oracle, X, rows, y, None, abi.Params.make(nbins=64, max_depth=5))
    assert len(levels) == 5


def test_sparse_storage_rejects_weights(lib):
    from h2ok_hist import capi

    h = C.c_void_p()
    lib.call("h2ok_hist_create", 0, C.byref(h))
    try:
        with pytest.raises(capi.H2okError) as e:
            lib.call("h2ok_hist_tree_create", h, 1000, 8, abi.TREE_SPARSE | abi.TREE_WEIGHTED)
        assert e.value.code == abi.ERR_UNSUPPORTED
        with pytest.raises(capi.H2okError):
            lib.call("h2ok_hist_tree_create", h, 1000, 8, 8)
    finally:
        lib.call("h2ok_hist_destroy", h)



# ------------------------------------------------------------------------------------------------ categorical columns
def categorical_case(oracle, rows, cols, seed, na):
    """Columns 1, 3 and 5 become categorical (codes 0..k-1, k = 5, 37, 300), and the target depends on membership in
    scattered sets of categories, so the winning splits are bit-set splits; column 4 is an integer column."""
    X, y, _ = oracle.synth(abi.SHAPE_WIDE64, seed, 0, rows, cols, na_permille=na)
    rng = np.random.default_rng(seed)
    is_int = [0] * cols
    yy = y[:rows].astype(np.float64) * 0.2
    for c, k in ((1, 5), (3, 37), (5, 300)):
        if c >= cols:
            continue
        codes = rng.integers(0, k, rows).astype(np.float32)
        nan = np.isnan(X[c, :rows])
        X[c, :rows] = np.where(nan, np.nan, codes)
        member = rng.random(k) < 0.5
        yy += np.where(nan, 0.0, member[codes.astype(np.int64)] * (1.0 + c))
        is_int[c] = 2
    if cols > 4:
        X[4, :rows] = np.floor(X[4, :rows] * 3)
        is_int[4] = 1
    y[:rows] = yy.astype(np.float32)
    return X, y, is_int


@pytest.mark.parametrize("rows,cols,nbins,depth,na", [(4099, 6, 64, 5, 0), (60001, 9, 512, 6, 20), (30011, 40, 1024, 5, 5)])
def test_loop_categorical_columns_match_oracle(eshard, oracle, rows, cols, nbins, depth, na):
    """is_int == 2: the split search sorts the categories by mean response, the split carries a bit set, rows are routed
    by their category, children keep the parent's range of the split column — every level against the oracle."""
    X, y, is_int = categorical_case(oracle, rows, cols, 900 + rows, na)
    params = abi.Params.make(nbins=nbins, max_depth=depth, min_rows=5)
    levels = check_tree(eshard, oracle, X, rows, y, None, params, is_int=is_int, host_out=True)
    assert any((lv["splits"]["equal"] == 2).any() for lv in levels)


def test_sparse_storage_categorical_columns(sshard, oracle):
    rows, cols = 30011, 70
    X, y, is_int = categorical_case(oracle, rows, cols, 77, 10)
    X[6:, :rows] = np.where(np.arange(rows)[None, :] % 7 < 5, 0.0, X[6:, :rows])  # the other columns mostly zero
    levels = check_tree(sshard, oracle, X, rows, y, None, abi.Params.make(nbins=512, max_depth=5, min_rows=5), is_int=is_int)
    assert any((lv["splits"]["equal"] == 2).any() for lv in levels)

# ------------------------------------------------------------------------------------------------ u8 storage
@pytest.fixture(scope="module")
def ushard():
    from h2ok_hist.driver import EngineShard

    s = EngineShard(0, u8=True)
    yield s
    s.close()


def codes_of(oracle, X, rows, nbins):
    """The oracle's pre-binning: every column's cells against its root descriptor; NA -> NaN, codes as floats."""
    root = oracle.root_descs(X, rows, abi.Params.make(nbins=nbins))
    codes = np.full(X.shape, np.nan, np.float32)
    for c in range(X.shape[0]):
        if root[c]["nbin"] > 0:
            b = oracle.bin_many(X[c, :rows], root[c]).astype(np.float32)
            b[b == root[c]["nbin"]] = np.nan  # the NA bucket
            codes[c, :rows] = b
    return root, codes


def test_prebin_matches_the_oracles_bins(ushard, oracle):
    rows, cols = 70001, 9
    X, _, _ = oracle.synth(abi.SHAPE_WIDE64, 11, 0, rows, cols, na_permille=20)
    X[3, :rows] = 2.5  # a constant column: untracked, all NA codes
    root, codes = codes_of(oracle, X, rows, 255)
    dX = torch.from_numpy(X).cuda()
    out = torch.full((cols, X.shape[1]), 7, dtype=torch.uint8, device="cuda")
    ushard.lib.call("h2ok_hist_prebin_dev", ushard.ctx, C.c_void_p(dX.data_ptr()), X.shape[1], rows, cols, abi.ptr(root),
                    C.c_void_p(out.data_ptr()), X.shape[1])
    got = out[:, :rows].cpu().numpy()
    want = np.where(np.isnan(codes[:, :rows]), 255, codes[:, :rows]).astype(np.uint8)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("rows,cols,depth,na,nbins_pre", [(7, 3, 3, 0, 255), (30011, 70, 6, 10, 255), (60001, 28, 7, 0, 200),
                                                           (20000, 33, 5, 30, 64)])
def test_u8_storage_matches_oracle_on_the_codes(ushard, oracle, rows, cols, depth, na, nbins_pre):
    """H2OK_TREE_U8: trees on pre-binned codes equal the oracle run on the code matrix as integer columns, bit for bit —
    histograms, ranges, splits, children, routing — whatever number of bins the codes were made with."""
    X, y, _ = oracle.synth(abi.SHAPE_WIDE64, 40 + rows, 0, rows, cols, na_permille=na)
    _, codes = codes_of(oracle, X, rows, nbins_pre)
    params = abi.Params.make(nbins=255, max_depth=depth, min_rows=1 if rows < 1000 else 5)
    check_tree(ushard, oracle, codes, rows, y, None, params, is_int=[1] * cols, host_out=(rows % 2 == 1))
    assert ushard.stats().range_errors == 0



def test_u8_storage_categorical_columns(ushard, oracle):
    """Codes of a categorical column are category ids: the same bit-set splits on the byte-wide storage."""
    rows, cols = 30011, 8
    X, y, _ = oracle.synth(abi.SHAPE_WIDE64, 55, 0, rows, cols, na_permille=10)
    _, codes = codes_of(oracle, X, rows, 200)
    rng = np.random.default_rng(3)
    cat = rng.integers(0, 150, rows)
    member = rng.random(150) < 0.5
    codes[2, :rows] = cat
    y[:rows] = (y[:rows] * 0.1 + member[cat] * 2.0).astype(np.float32)
    is_int = [1] * cols
    is_int[2] = 2
    levels = check_tree(ushard, oracle, codes, rows, y, None, abi.Params.make(nbins=255, max_depth=5, min_rows=5), is_int=is_int)
    assert any((lv["splits"]["equal"] == 2).any() for lv in levels)

def test_u8_storage_refuses_what_it_cannot_hold(ushard, oracle, lib):
    from h2ok_hist import capi
    from h2ok_hist.driver import TreeLoop

    rows, cols = 5000, 4
    X, y, _ = oracle.synth(abi.SHAPE_HIGGS28, 3, 0, rows, cols)
    _, codes = codes_of(oracle, X, rows, 255)
    ushard.set_data_tensors(torch.from_numpy(codes[:, :rows].copy()), torch.from_numpy(y[:rows].copy()))
    with pytest.raises(capi.H2okError) as e:  # fewer bins than codes: step-1 bins are not guaranteed
        TreeLoop(ushard, abi.Params.make(nbins=64, max_depth=3))
    assert e.value.code == abi.ERR_UNSUPPORTED
    with pytest.raises(capi.H2okError) as e:  # codes are integer columns
        TreeLoop(ushard, abi.Params.make(nbins=255, max_depth=3), is_int=[0] * cols)
    assert e.value.code == abi.ERR_ARG
    ushard.set_data_tensors(torch.from_numpy(X[:, :rows].copy()), torch.from_numpy(y[:rows].copy()))  # floats, not codes
    loop = TreeLoop(ushard, abi.Params.make(nbins=255, max_depth=3))
    loop.begin_tree()
    with pytest.raises(capi.H2okError) as e:
        loop.level()
    assert e.value.code == abi.ERR_ARG and "code" in str(e.value)
    h = C.c_void_p()
    lib.call("h2ok_hist_create", 0, C.byref(h))
    try:
        with pytest.raises(capi.H2okError):
            lib.call("h2ok_hist_tree_create", h, 100, 4, abi.TREE_U8 | abi.TREE_SPARSE)
        lib.call("h2ok_hist_tree_create", h, 100, 4, 0)
        buf = torch.zeros(4096, dtype=torch.uint8, device="cuda")
        with pytest.raises(capi.H2okError) as e:  # bins into a dense shard
            lib.call("h2ok_hist_tree_load_bins_dev", h, 0, 4, C.c_void_p(buf.data_ptr()), 100)
        assert e.value.code == abi.ERR_STATE
    finally:
        lib.call("h2ok_hist_destroy", h)


def test_loop_rejects_wrong_call_order(lib):
    h = C.c_void_p()
    lib.call("h2ok_hist_create", 0, C.byref(h))
    from h2ok_hist import capi

    try:
        p = abi.Params.make(nbins=64, max_depth=3)
        with pytest.raises(capi.H2okError) as e:
            lib.call("h2ok_hist_loop_create", h, C.byref(p), None)
        assert e.value.code == abi.ERR_STATE
        with pytest.raises(capi.H2okError) as e:
            lib.call("h2ok_hist_loop_level", h, None, None, None)
        assert e.value.code == abi.ERR_STATE
        lib.call("h2ok_hist_tree_create", h, 1000, 3, 0)
        lib.call("h2ok_hist_tree_load_synth", h, abi.SHAPE_HIGGS28, 7, 0, 0)
        lib.call("h2ok_hist_loop_create", h, C.byref(p), None)
        with pytest.raises(capi.H2okError) as e:  # no tree begun
            lib.call("h2ok_hist_loop_level", h, None, None, None)
        assert e.value.code == abi.ERR_STATE
        lib.call("h2ok_hist_loop_begin_tree_synth", h, 7, 0, 0)
        done, nch = C.c_int32(), C.c_int32()
        for _ in range(3):
            lib.call("h2ok_hist_loop_level", h, None, C.byref(done), C.byref(nch))
            if nch.value == 0:
                break
        with pytest.raises(capi.H2okError) as e:  # the tree is finished
            lib.call("h2ok_hist_loop_level", h, None, None, None)
        assert e.value.code == abi.ERR_STATE
        bad = abi.Params.make(nbins=5000, max_depth=3)
        with pytest.raises(capi.H2okError) as e:
            lib.call("h2ok_hist_loop_create", h, C.byref(bad), None)
        assert e.value.code == abi.ERR_UNSUPPORTED
    finally:
        lib.call("h2ok_hist_destroy", h)


def test_loop_synth_entry_points_need_no_torch_memory(lib, oracle):
    """tree_load_synth + loop_begin_tree_synth (what a torch-free caller such as h2ok-bench uses) give the same tree as
    the oracle on the generator's own data."""
    rows, cols = 70001, 28
    h = C.c_void_p()
    lib.call("h2ok_hist_create", 0, C.byref(h))
    try:
        p = abi.Params.make(nbins=64, max_depth=5)
        lib.call("h2ok_hist_tree_create", h, rows, cols, 0)
        lib.call("h2ok_hist_tree_load_synth", h, abi.SHAPE_HIGGS28, 20261106, 0, 0)
        lib.call("h2ok_hist_loop_create", h, C.byref(p), None)
        lib.call("h2ok_hist_loop_set_options", h, 1, 0)
        lib.call("h2ok_hist_loop_begin_tree_synth", h, 20261106, 0, 0)
        nch = C.c_int32(1)
        while nch.value:
            lib.call("h2ok_hist_loop_level", h, None, None, C.byref(nch))
        got = (C.c_uint64 * 3)()
        lib.call("h2ok_hist_loop_checksum", h, 0, got)
        X, y, _ = oracle.synth(abi.SHAPE_HIGGS28, 20261106, 0, rows, cols)
        levels, _, leaf, _ = oracle_lib.train_tree_q(oracle, X, rows, y, None, p)
        assert [int(x) for x in got] == oracle_lib.tree_checksum(levels, leaf)
    finally:
        lib.call("h2ok_hist_destroy", h)


# ------------------------------------------------------------------------------------------------ a9 on the loop
@pytest.mark.parametrize("sparse", [False, True])
def test_multinomial_k10_on_the_loop_tracks_oracle(eshard, sshard, oracle, sparse):
    """BASELINE config 4's trainer: 10-class multinomial GBM (ten trees per iteration on softmax residuals) through the
    level loop inside the library, on dense and on sparse storage.  The first iteration starts from zero scores (exp is
    exact on both sides): its ten trees and leaf values equal the oracle's bit for bit; later iterations go through
    expf (device vs libm), so the check is the logloss trace to 1e-6."""
    from h2ok_hist.driver import TreeLoop
    from h2ok_hist.gbm import GBMTrainer

    sh = sshard if sparse else eshard
    rows, cols, K = 30000, 70, 10
    X, y, _ = oracle.synth(abi.SHAPE_SPARSE, 5, 0, rows, cols, na_permille=0 if sparse else 5)
    score = X[:K, :rows].T @ np.linspace(1.0, 2.0, K * K).reshape(K, K).astype(np.float32) + y[:rows, None] * 0.2
    label = score.argmax(axis=1).astype(np.float32)
    params = abi.Params.make(nbins=32, max_depth=4, min_rows=5)
    sh.set_data_tensors(torch.from_numpy(X[:, :rows].copy()), torch.from_numpy(label))
    loop = TreeLoop(sh, params)
    tr = GBMTrainer(loop, torch.from_numpy(label).cuda(), abi.DIST_MULTINOMIAL, learn_rate=0.3, n_classes=K)
    assert tr.init_f() == 0.0
    ref = oracle_lib.OracleGBM(oracle, X, rows, label, None, params, abi.DIST_MULTINOMIAL, 0.3, n_classes=K)

    def logloss(f):
        z = f.astype(np.float64)
        z = z - z.max(axis=0, keepdims=True)
        logp = z - np.log(np.exp(z).sum(axis=0, keepdims=True))
        return float(-logp[label.astype(np.int64), np.arange(rows)].mean())

    trace = [tr.loss()]
    assert trace[0] == pytest.approx(np.log(K), rel=1e-12)
    for it in range(2):
        gs = tr.step()
        grs = ref.step()
        torch.cuda.synchronize()
        if it == 0:
            for k in range(K):
                assert gs[k].tobytes() == grs[k].tobytes(), f"class {k}: leaf values differ"
                for (s_gpu, ids_gpu), (s_ref, ids_ref) in zip(tr.trees[k]["levels"], ref.trees[k]["levels"]):
                    assert s_gpu.tobytes() == s_ref.tobytes() and np.array_equal(ids_gpu, ids_ref)
            assert np.array_equal(tr.f[:, :rows].cpu().numpy(), ref.f)
        trace.append(tr.loss())
        assert trace[-1] == pytest.approx(logloss(ref.f), rel=1e-6)
    assert all(b < a for a, b in zip(trace, trace[1:]))
    assert len(tr.trees) == 2 * K


def test_forest_equals_the_trees_built_one_by_one(sshard, oracle):
    """h2ok_hist_loop_create_forest: K trees in lockstep on the sparse storage are, tree by tree, the trees the loop
    builds alone — same splits at every node, same leaf for every row (up to the numbering of the nodes, which is flat
    over the forest) — and hence the same result checksum parts per tree."""
    from h2ok_hist.driver import TreeLoop

    rows, cols, K = 40001, 130, 5
    X, y, _ = oracle.synth(abi.SHAPE_SPARSE, 77, 0, rows, cols, na_permille=5)
    ys = np.stack([np.roll(y[:rows], 17 * k) * (1.0 + 0.25 * k) + 0.1 * k * X[k % cols, :rows] for k in range(K)]).astype(np.float32)
    params = abi.Params.make(nbins=32, max_depth=5, min_rows=5)
    sshard.set_data_tensors(torch.from_numpy(X[:, :rows].copy()), torch.from_numpy(ys[0].copy()))
    ld = (rows + 63) // 64 * 64
    dys = torch.zeros((K, ld), dtype=torch.float32, device="cuda")
    dys[:, :rows] = torch.from_numpy(ys)
    alone = []
    loop = TreeLoop(sshard, params)
    for k in range(K):  # the trees one by one: per level the split records, at the end every row's leaf structure
        loop.begin_tree(y=dys[k])
        levels = []
        while loop.n_nodes_left() > 0:
            loop.level()
            lv = loop.done_level()
            levels.append((lv["splits"].copy(), lv["child_ids"].copy()))
        alone.append((levels, sshard.leaf[:rows].cpu().numpy().copy()))
    forest = TreeLoop(sshard, params, forest=K)
    forest.begin_forest(dys)
    flevels = []
    while forest.n_nodes_left() > 0:
        forest.level()
        lv = forest.done_level()
        flevels.append((lv["splits"].copy(), lv["child_ids"].copy(), lv["n"]))
    v = forest.view()
    assert v.n_trees == K
    leaf_all = forest._dev_array(v.d_leaf, 4 * K * rows).view(np.int32).reshape(K, rows)
    # walk the forest's flat levels class by class: the flat node list of a level is tree 0's nodes, then tree 1's, ...
    counts = [1] * K
    for d, (fs, fids, n) in enumerate(flevels):
        assert n == sum(counts)
        off, nxt = 0, []
        for k in range(K):
            if counts[k] == 0:
                nxt.append(0)
                continue
            s_alone, ids_alone = alone[k][0][d]
            mine = fs[off:off + counts[k]]
            for name in ("col", "bin", "na_dir", "splat", "se_before", "se_left", "se_right", "n_left", "n_right",
                         "pred_left", "pred_right"):
                assert np.array_equal(mine[name], s_alone[name]), (d, k, name)
            # same children exist (ids differ by the forest's flat numbering)
            assert np.array_equal(fids[2 * off:2 * (off + counts[k])] >= 0, ids_alone >= 0)
            nxt.append(int((ids_alone >= 0).sum()))
            off += counts[k]
        counts = nxt
    # leaves: the partition of the rows must be the same (leaf codes are numbered differently)
    for k in range(K):
        a, b = alone[k][1], leaf_all[k]
        assert (a >= 0).all() and (b >= 0).all()
        pairs = np.unique(np.stack([a, b]), axis=1)
        assert len(np.unique(pairs[0])) == pairs.shape[1] == len(np.unique(pairs[1]))  # a bijection between the codes


def test_multinomial_k10_as_a_forest_tracks_oracle(sshard, oracle):
    """The K = 10 trainer on the forest path: first iteration's scores bit-identical to the oracle's, logloss after."""
    from h2ok_hist.driver import TreeLoop
    from h2ok_hist.gbm import GBMTrainer

    rows, cols, K = 30000, 70, 10
    X, y, _ = oracle.synth(abi.SHAPE_SPARSE, 5, 0, rows, cols)
    score = X[:K, :rows].T @ np.linspace(1.0, 2.0, K * K).reshape(K, K).astype(np.float32) + y[:rows, None] * 0.2
    label = score.argmax(axis=1).astype(np.float32)
    params = abi.Params.make(nbins=32, max_depth=4, min_rows=5)
    sshard.set_data_tensors(torch.from_numpy(X[:, :rows].copy()), torch.from_numpy(label))
    loop = TreeLoop(sshard, params, forest=K)
    tr = GBMTrainer(loop, torch.from_numpy(label).cuda(), abi.DIST_MULTINOMIAL, learn_rate=0.3, n_classes=K)
    tr.init_f()
    ref = oracle_lib.OracleGBM(oracle, X, rows, label, None, params, abi.DIST_MULTINOMIAL, 0.3, n_classes=K)

    def logloss(f):
        z = f.astype(np.float64)
        z = z - z.max(axis=0, keepdims=True)
        logp = z - np.log(np.exp(z).sum(axis=0, keepdims=True))
        return float(-logp[label.astype(np.int64), np.arange(rows)].mean())

    trace = [tr.loss()]
    for it in range(2):
        tr.step()
        ref.step()
        torch.cuda.synchronize()
        if it == 0:
            assert np.array_equal(tr.f[:, :rows].cpu().numpy(), ref.f)
        trace.append(tr.loss())
        assert trace[-1] == pytest.approx(logloss(ref.f), rel=1e-6)
    assert all(b < a for a, b in zip(trace, trace[1:]))


def test_gaussian_on_the_loop_is_bit_exact(eshard, oracle):
    from h2ok_hist.driver import TreeLoop
    from h2ok_hist.gbm import GBMTrainer

    rows, cols = 30011, 10
    X, y, w = oracle.synth(abi.SHAPE_WIDE64, 321, 0, rows, cols, na_permille=10, weight_mode=1)
    label = (y * 3 + 1).astype(np.float32)
    params = abi.Params.make(nbins=64, max_depth=4, min_rows=5)
    eshard.set_data_tensors(torch.from_numpy(X[:, :rows].copy()), torch.from_numpy(label), torch.from_numpy(w[:rows].copy()))
    loop = TreeLoop(eshard, params)
    tr = GBMTrainer(loop, torch.from_numpy(label).cuda(), abi.DIST_GAUSSIAN, learn_rate=0.3)
    f0 = tr.init_f()
    ref = oracle_lib.OracleGBM(oracle, X, rows, label, w[:rows].copy(), params, abi.DIST_GAUSSIAN, 0.3)
    ref.init_f(f0)
    for t in range(3):
        g, gr = tr.step(), ref.step()
        torch.cuda.synchronize()
        assert g.tobytes() == gr.tobytes(), f"tree {t}: leaf values differ"
        assert np.array_equal(tr.f.cpu().numpy(), ref.f), f"tree {t}: f differs"
