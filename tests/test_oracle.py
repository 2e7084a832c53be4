"""CPU tests of the oracle: hand-derived known answers for every H2O-3 rule it restates (SURVEY.md §8(c):
the reference pins nothing for this path, so these paper-computed cases are the pins), an independent numpy
restatement, and sklearn's HistogramBuilder as a second opinion on accumulation."""
import math

import numpy as np
import pytest

from h2ok_hist import abi


def desc(mn, maxex, step, nbin, is_int=0):
    d = np.zeros(1, abi.COL_DESC)
    d["min"], d["maxex"], d["step"], d["nbin"], d["is_int"] = mn, maxex, step, nbin, is_int
    return d[0]


# ---------------------------------------------------------------- generator
def test_splitmix64_known_vectors(oracle):
    # Vigna's splitmix64 reference stream: state 0 → 0xE220A8397B1DCDAF; state 1234567 → 6457827717110365317
    assert oracle.lib.h2ok_oracle_splitmix64(0) == 0xE220A8397B1DCDAF
    assert oracle.lib.h2ok_oracle_splitmix64(1234567) == 6457827717110365317


def test_generator_shapes(oracle):
    X, y, w = oracle.synth(abi.SHAPE_HIGGS28, 20261106, 0, 4000, 28)
    assert X.shape == (28, 4000) and not np.isnan(X).any()
    for c in (8, 12, 16, 20):  # b-tag-like columns take exactly three values
        assert set(np.unique(X[c]).tolist()) == {0.0, np.float32(1.0865), np.float32(2.173)}
    assert X[1].min() >= -2.5 and X[1].max() < 2.5
    assert X[0].min() >= 0.25 and X[0].max() < 4.25
    assert X[21].min() >= 0.3 and X[21].max() < 3.3
    # row-offset invariance: a shard generated at row0 equals the slice of the whole
    Xs, ys, _ = oracle.synth(abi.SHAPE_HIGGS28, 20261106, 1000, 500, 28)
    assert np.array_equal(Xs[:, :500], X[:, 1000:1500]) and np.array_equal(ys, y[1000:1500])
    Xn, _, wn = oracle.synth(abi.SHAPE_WIDE64, 7, 0, 20000, 64, na_permille=10, weight_mode=1)
    frac = np.isnan(Xn).mean()
    assert 0.005 < frac < 0.015
    assert 0.01 < (wn == 0).mean() < 0.03 and wn.max() < 1.5 and wn[wn > 0].min() >= 0.5
    Xz, _, _ = oracle.synth(abi.SHAPE_SPARSE, 7, 0, 20000, 8)
    assert 0.88 < (Xz == 0).mean() < 0.92


# ---------------------------------------------------------------- a1 descriptors
def test_find_maxex(oracle):
    assert oracle.find_maxex(3.0, 1) == 4.0  # integer column: +1
    assert oracle.find_maxex(3.0, 0) == np.nextafter(3.0, np.inf)  # real: +ulp (double)
    assert oracle.find_maxex(0.0, 0) == 5e-324  # Math.ulp(0.0) is Double.MIN_VALUE
    assert oracle.find_maxex(-2.5, 0) == -2.5 + (np.nextafter(2.5, np.inf) - 2.5)
    assert oracle.find_maxex(1e300, 1) == 1e300 + (np.nextafter(1e300, np.inf) - 1e300)  # ulp > 1 wins


def test_make_desc_rules(oracle):
    d, ok = oracle.make_desc(0.0, 10.0, 0, 4)
    assert ok and d["nbin"] == 4 and d["step"] == 0.4
    d, ok = oracle.make_desc(0.0, 4.0, 1, 256)  # integer column with maxEx-min <= nbins → one bin per integer
    assert ok and d["nbin"] == 4 and d["step"] == 1.0
    d, ok = oracle.make_desc(0.0, 1000.0, 1, 256)  # integer column with a wide range keeps nbins
    assert ok and d["nbin"] == 256 and d["step"] == 0.256
    d, ok = oracle.make_desc(2.0, 2.0, 0, 256)  # constant → no histogram
    assert not ok and d["nbin"] == 0
    d, ok = oracle.make_desc(-7.0, -3.0, 1, 256)
    assert ok and d["nbin"] == 4 and d["step"] == 1.0


def test_root_descs(oracle):
    rows = 64
    X = np.zeros((4, rows), np.float32)
    X[0] = np.linspace(-1, 1, rows)
    X[1] = 5.0  # constant → untracked
    X[2] = np.arange(rows) % 3  # integer valued
    X[3] = np.nan  # all NA → untracked
    p = abi.Params.make(nbins=16)
    d = oracle.root_descs(X, rows, p, is_int=[0, 0, 1, 0])
    assert d["nbin"].tolist() == [16, 0, 3, 0]
    assert d[0]["min"] == -1.0 and d[0]["maxex"] == np.nextafter(1.0, np.inf)
    assert d[0]["step"] == 16 / (d[0]["maxex"] - d[0]["min"])
    assert d[2]["min"] == 0.0 and d[2]["maxex"] == 3.0 and d[2]["step"] == 1.0


# ---------------------------------------------------------------- a2 bin
def test_bin_known_answers(oracle):
    d = desc(0.0, 10.0, 0.4, 4)
    cases = {0.0: 0, 2.49: 0, 2.5: 1, 4.99: 1, 5.0: 2, 7.5: 3, 9.99: 3,
             10.0: 3,  # pos == nbin → round-off clamp (idx--)
             -1.0: 0, 25.0: 3,  # outside the range: clamped to the end bins
             float("-inf"): 0, float("inf"): 3, float("nan"): 4}  # NA bucket is slot nbin
    for x, b in cases.items():
        assert oracle.bin(x, d) == b, x


def test_bin_is_fp64_not_fp32(oracle):
    # fp64 arithmetic decides boundary cases an fp32 shortcut gets wrong
    mn, maxex = 0.1, 0.1 + 0.7
    step = 256 / (maxex - mn)
    d = desc(mn, maxex, step, 256)
    xs = (np.float32(0.1) + np.arange(0, 200000, dtype=np.float32) * np.float32(3.5e-6)).astype(np.float32)
    got = oracle.bin_many(xs, d)
    want = np.clip(((xs.astype(np.float64) - mn) * step).astype(np.int64), 0, 255)
    assert np.array_equal(got, want)
    f32 = np.clip(((xs - np.float32(mn)) * np.float32(step)).astype(np.int64), 0, 255)
    assert (f32 != want).any(), "case should separate fp32 from fp64 binning"


def test_bin_value_just_below_maxex(oracle):
    mx = np.float32(3.25)
    maxex = oracle.find_maxex(float(mx), 0)
    d, ok = oracle.make_desc(1.0, maxex, 0, 256)
    assert ok
    assert oracle.bin(float(mx), d) == 255
    assert oracle.bin(1.0, d) == 0


# ---------------------------------------------------------------- a3 updateHisto
def test_build_level_paper_case(oracle):
    # one node, one column, 4 bins of width 1; bins: 0,1,1,3,NA
    X = np.array([[0.5, 1.5, 1.7, 3.2, np.nan, 0, 0, 0]], np.float32)
    y = np.array([1, 2, -1, 0.5, 4], np.float32)
    nid = np.zeros(5, np.int32)
    d = np.array([[desc(0.0, 4.0, 1.0, 4)]], abi.COL_DESC)
    hist, absy, min2, maxin = oracle.build_level(X, 5, y, None, nid, d, 4)
    want = np.zeros((5, 3))
    want[0] = [1, 1, 1]
    want[1] = [2, 1, 5]
    want[3] = [1, 0.5, 0.25]
    want[4] = [1, 4, 16]
    assert np.array_equal(hist[0, 0], want)
    assert absy[0, 0].tolist() == [1, 3, 0, 0.5, 4]
    assert min2[0, 0] == np.float32(0.5) and maxin[0, 0] == np.float32(3.2)  # NaN never updates min/max


def test_build_level_weights_and_nodes(oracle):
    X = np.array([[0.5, 1.5, 1.7, 3.2], [10, 20, 30, 40]], np.float32)
    y = np.array([1, 2, -1, 0.5], np.float32)
    w = np.array([2, 0, 0.5, 1], np.float32)  # zero-weight row is skipped entirely (DRF rule in updateHisto)
    nid = np.array([0, 0, 1, -1], np.int32)  # last row already sits in a leaf
    d = np.zeros((2, 2), abi.COL_DESC)
    d[:, 0] = desc(0.0, 4.0, 1.0, 4)
    d[:, 1] = desc(10.0, 50.0, 0.1, 4)
    d[1, 1]["nbin"] = 0  # column 1 untracked in node 1
    hist, _, min2, maxin = oracle.build_level(X, 4, y, w, nid, d, 4)
    assert hist[0, 0, 0].tolist() == [2, 2, 2] and hist[0, 0].sum() == 6
    assert hist[1, 0, 1].tolist() == [0.5, -0.5, 0.5] and hist[1, 0].sum() == 0.5
    assert hist[0, 1, 0].tolist() == [2, 2, 2]
    assert hist[1, 1].sum() == 0
    assert min2[0, 0] == 0.5 and maxin[0, 0] == 0.5  # the zero-weight row (1.5) is not observed
    assert np.isinf(min2[1, 1]) and np.isinf(maxin[1, 1])


def numpy_build_level(X, rows, y, w, nid, descs, max_bins):
    """Independent restatement (np.add.at) — catches oracle-and-kernel-share-a-bug."""
    cols = X.shape[0]
    n_nodes = descs.shape[0]
    hist = np.zeros((n_nodes, cols, max_bins + 1, 3))
    wd = np.ones(rows) if w is None else w[:rows].astype(np.float64)
    yd = y[:rows].astype(np.float64)
    for c in range(cols):
        x = X[c, :rows].astype(np.float64)
        for g in range(n_nodes):
            d = descs[g, c]
            if d["nbin"] <= 0:
                continue
            sel = (nid[:rows] == g) & (wd != 0)
            xs = x[sel]
            with np.errstate(invalid="ignore", over="ignore"):
                pos = (xs - d["min"]) * d["step"]
                b = np.where(pos < 0, 0, np.where(pos >= d["nbin"], d["nbin"] - 1, np.nan_to_num(pos, nan=0)).astype(np.int64))
            b = np.where(np.isnan(xs), d["nbin"], b)
            wy = wd[sel] * yd[sel]
            np.add.at(hist[g, c, :, 0], b, wd[sel])
            np.add.at(hist[g, c, :, 1], b, wy)
            np.add.at(hist[g, c, :, 2], b, wy * yd[sel])
    return hist


@pytest.mark.parametrize("weight_mode,na", [(0, 0), (1, 20)])
def test_build_level_vs_numpy(oracle, weight_mode, na):
    rows, cols, nb = 5000, 6, 32
    X, y, w = oracle.synth(abi.SHAPE_WIDE64, 11, 0, rows, cols, na_permille=na, weight_mode=weight_mode)
    rng = np.random.default_rng(0)
    nid = rng.integers(-1, 3, rows).astype(np.int32)
    root = oracle.root_descs(X, rows, abi.Params.make(nbins=nb))
    descs = np.stack([root, root, root])
    descs[1]["min"] += 0.25  # different ranges per node, some values fall outside and clamp
    descs[2, 3]["nbin"] = 0
    hist, _, _, _ = oracle.build_level(X, rows, y, w, nid, descs, nb)
    ref = numpy_build_level(X, rows, y, w, nid, descs, nb)
    assert np.array_equal(hist[..., 0], ref[..., 0])
    assert np.allclose(hist, ref, rtol=1e-12, atol=1e-12)
    hist_mt, _, _, _ = oracle.build_level(X, rows, y, w, nid, descs, nb, threads=4)
    assert np.array_equal(hist_mt, hist)  # column-parallel run is the same arithmetic


def test_accumulation_vs_sklearn(oracle):
    hb = pytest.importorskip("sklearn.ensemble._hist_gradient_boosting.histogram")
    rows, cols, nb = 4000, 5, 64
    X, y, _ = oracle.synth(abi.SHAPE_HIGGS28, 3, 0, rows, cols)
    root = oracle.root_descs(X, rows, abi.Params.make(nbins=nb))
    binned = np.stack([oracle.bin_many(X[c, :rows], root[c]) for c in range(cols)], axis=1).astype(np.uint8)
    binned = np.asfortranarray(binned)
    g = y.astype(np.float32)
    h = (y * y).astype(np.float32)
    builder = hb.HistogramBuilder(binned, nb + 1, g, h, hessians_are_constant=False, n_threads=1)
    sk = np.asarray(builder.compute_histograms_brute(np.arange(rows, dtype=np.uint32)))
    hist, _, _, _ = oracle.build_level(X, rows, y, None, np.zeros(rows, np.int32), root.reshape(1, cols), nb)
    assert np.array_equal(sk["count"][:, :nb].astype(np.float64), hist[0, :, :nb, 0])
    assert np.allclose(sk["sum_gradients"][:, :nb], hist[0, :, :nb, 1], rtol=1e-9, atol=1e-9)
    assert np.allclose(sk["sum_hessians"][:, :nb], hist[0, :, :nb, 2], rtol=1e-6, atol=1e-9)


# ---------------------------------------------------------------- fixed point
def test_quantisation_contract(oracle):
    rows = 3000
    _, y, w = oracle.synth(abi.SHAPE_HIGGS28, 5, 0, rows, 28, weight_mode=1)
    q, m = oracle.choose_quant(y, w, 25)
    for s, mx in ((q.scale_w, m[0]), (q.scale_wy, m[1]), (q.scale_wyy, m[2])):
        assert math.frexp(s)[0] == 0.5  # power of two
        assert 2.0**24 <= mx * s <= 2.0**25
    wq, wyq, wyyq = oracle.quantize_rows(y, w, q)
    wy = w.astype(np.float64) * y.astype(np.float64)
    assert np.array_equal(wyq, np.rint(wy * q.scale_wy).astype(np.int32))
    assert np.array_equal(wyyq, np.rint(wy * y.astype(np.float64) * q.scale_wyy).astype(np.int32))
    assert np.array_equal(wq == 0, w == 0)
    q0, _ = oracle.choose_quant(y, None, 25)
    assert q0.scale_w == 1.0 and not q0.weighted


@pytest.mark.parametrize("weight_mode", [0, 1])
def test_quantised_level_matches_fp64_level(oracle, weight_mode):
    rows, cols, nb = 6000, 7, 64
    X, y, w = oracle.synth(abi.SHAPE_WIDE64, 17, 0, rows, cols, na_permille=5, weight_mode=weight_mode)
    nid = (np.arange(rows) % 3).astype(np.int32)
    root = oracle.root_descs(X, rows, abi.Params.make(nbins=nb))
    descs = np.stack([root] * 3)
    q, m = oracle.choose_quant(y, w, 25)
    wq, wyq, wyyq = oracle.quantize_rows(y, w, q)
    acc, min2q, maxinq = oracle.build_level_q(X, rows, nid, wq, wyq, wyyq, descs, nb)
    hq = oracle.finalize(acc, q, nb)
    hist, absy, min2, maxin = oracle.build_level(X, rows, y, w, nid, descs, nb)
    assert np.array_equal(min2q, min2) and np.array_equal(maxinq, maxin)
    if not weight_mode:
        assert np.array_equal(hq[..., 0], hist[..., 0])  # counts are exact
    cnt = np.maximum(hist[..., 0] if not weight_mode else hq[..., 0] * 0 + 1, 1)
    # every channel: |err| <= rows_in_bin * half a quantum, far inside north_star's 1e-5 relative
    n_in_bin = oracle.build_level(X, rows, np.ones(rows, np.float32), None, nid, descs, nb)[0][..., 0]
    assert (np.abs(hq[..., 1] - hist[..., 1]) <= n_in_bin * 0.5 / q.scale_wy + 1e-300).all()
    assert (np.abs(hq[..., 2] - hist[..., 2]) <= n_in_bin * 0.5 / q.scale_wyy + 1e-300).all()
    assert (np.abs(hq[..., 1] - hist[..., 1]) <= 1e-5 * absy + n_in_bin * 0.5 / q.scale_wy).all()
    del cnt


def test_shard_then_sum_is_exact_in_fixed_point(oracle):
    rows, cols, nb = 5000, 4, 32
    X, y, _ = oracle.synth(abi.SHAPE_HIGGS28, 23, 0, rows, cols)
    nid = (np.arange(rows) % 2).astype(np.int32)
    root = oracle.root_descs(X, rows, abi.Params.make(nbins=nb))
    descs = np.stack([root] * 2)
    q, _ = oracle.choose_quant(y, None, 25)
    _, wyq, wyyq = oracle.quantize_rows(y, None, q)
    whole, mn, mx = oracle.build_level_q(X, rows, nid, None, wyq, wyyq, descs, nb)
    parts, mns, mxs = [], [], []
    for lo, hi in ((0, 1700), (1700, 1704), (1704, rows)):
        Xs = np.ascontiguousarray(np.pad(X[:, lo:hi], ((0, 0), (0, (-(hi - lo)) % 4))))
        a, m1, m2 = oracle.build_level_q(Xs, hi - lo, nid[lo:hi].copy(), None, wyq[lo:hi].copy(), wyyq[lo:hi].copy(), descs, nb)
        parts.append(a)
        mns.append(m1)
        mxs.append(m2)
    assert np.array_equal(sum(parts), whole)  # MRTask.reduce as an integer sum: identical for any sharding
    assert np.array_equal(np.minimum.reduce(mns), mn) and np.array_equal(np.maximum.reduce(mxs), mx)


# ---------------------------------------------------------------- a7 split search
def hist_from(counts, ysum, yysum, na=(0, 0, 0), nb=None):
    nb = nb or len(counts)
    h = np.zeros((1, 1, nb + 1, 3))
    h[0, 0, : len(counts), 0] = counts
    h[0, 0, : len(counts), 1] = ysum
    h[0, 0, : len(counts), 2] = yysum
    h[0, 0, nb] = na
    return h


def test_find_split_paper_case(oracle):
    # bins 0,1 hold y=0; bins 2,3 hold y=1 → perfect split at bin 2
    h = hist_from([10, 10, 10, 10], [0, 0, 10, 10], [0, 0, 10, 10])
    d = np.array([[desc(0.0, 8.0, 0.5, 4)]], abi.COL_DESC)
    s = oracle.find_splits(h, d, 10.0, 1e-5)[0]
    assert s["col"] == 0 and s["bin"] == 2
    assert s["se_before"] == 10.0 and s["se_left"] == 0 and s["se_right"] == 0  # 20 - 400/40
    assert s["n_left"] == 20 and s["n_right"] == 20 and s["pred_left"] == 0 and s["pred_right"] == 1
    assert s["na_dir"] == abi.NA_RIGHT  # no NAs seen: nLeft > nRight ? Left : Right
    assert s["splat"] == np.float32(4.0)  # binAt(2) = 0 + 2/0.5, both neighbours non-empty


def test_find_split_rules(oracle):
    d = np.array([[desc(0.0, 8.0, 0.5, 4)]], abi.COL_DESC)
    # min_rows: 15 rows needed each side → only b=2 qualifies even though b=1 is as good
    h = hist_from([10, 10, 10, 10], [0, 10, 10, 10], [0, 10, 10, 10])
    assert oracle.find_splits(h, d, 10.0, 0.0)[0]["bin"] == 1
    assert oracle.find_splits(h, d, 15.0, 0.0)[0]["bin"] == 2
    # tot < 2*min_rows → no split
    assert oracle.find_splits(h, d, 21.0, 0.0)[0]["col"] == -1
    # constant response → (float)var == 0 → no split
    hc = hist_from([10, 10, 10, 10], [10, 10, 10, 10], [10, 10, 10, 10])
    assert oracle.find_splits(hc, d, 1.0, 0.0)[0]["col"] == -1
    # min_split_improvement: gain 10 of 10 passes 1e-5; a tiny gain does not
    hn = hist_from([10, 10, 10, 10], [5, 5, 5, 5.001], [5, 5, 5, 5.001])
    assert oracle.find_splits(hn, d, 1.0, 0.5)[0]["col"] == -1
    # ties go to the bin nearest nbins>>1: symmetric response, b=1 and b=3 tie, b=2 is worse
    ht = hist_from([10, 10, 10, 10], [10, 0, 0, 10], [10, 0, 0, 10])
    s = oracle.find_splits(ht, d, 1.0, 0.0)[0]
    assert s["bin"] == 1  # |1-2| == |3-2|: first one found stays (strict <)
    # empty bins are skipped and splat is the midpoint of the gap
    he = hist_from([10, 0, 0, 10], [0, 0, 0, 10], [0, 0, 0, 10])
    s = oracle.find_splits(he, d, 1.0, 0.0)[0]
    assert s["bin"] == 3 and s["splat"] == np.float32((2.0 + 6.0) / 2)


def test_find_split_na_policies(oracle):
    d = np.array([[desc(0.0, 8.0, 0.5, 4)]], abi.COL_DESC)
    # NAs carry y=1 like the right side → best is a numeric split with NAs thrown right
    h = hist_from([10, 10, 10, 10], [0, 0, 10, 10], [0, 0, 10, 10], na=(10, 10, 10))
    s = oracle.find_splits(h, d, 5.0, 0.0)[0]
    assert s["bin"] == 2 and s["na_dir"] == abi.NA_RIGHT and s["n_right"] == 30 and s["se_left"] + s["se_right"] == 0
    # NAs carry y=0 → thrown left
    h = hist_from([10, 10, 10, 10], [0, 0, 10, 10], [0, 0, 10, 10], na=(10, 0, 0))
    s = oracle.find_splits(h, d, 5.0, 0.0)[0]
    assert s["bin"] == 2 and s["na_dir"] == abi.NA_LEFT and s["n_left"] == 30
    # response differs only between NA and non-NA → NA vs rest wins; routing threshold unused (-1)
    h = hist_from([10, 10, 10, 10], [0, 0, 0, 0], [0, 0, 0, 0], na=(20, 20, 20))
    s = oracle.find_splits(h, d, 5.0, 0.0)[0]
    assert s["na_dir"] == abi.NA_VS_REST and s["n_left"] == 40 and s["n_right"] == 20 and s["splat"] == -1.0
    assert s["se_before"] == pytest.approx(20 - 400 / 60)


def test_best_col_picks_min_se_first_on_ties(oracle):
    h = np.concatenate([hist_from([10, 10, 10, 10], [0, 0, 10, 10], [0, 0, 10, 10])] * 3, axis=1)
    h[0, 0, :4, 1] = [0, 5, 5, 10]  # column 0 is worse
    d = np.array([[desc(0.0, 8.0, 0.5, 4)] * 3], abi.COL_DESC)
    assert oracle.find_splits(h, d, 1.0, 0.0)[0]["col"] == 1
    d[0, 1]["nbin"] = 1  # nbins <= 1 is skipped by bestCol
    assert oracle.find_splits(h, d, 1.0, 0.0)[0]["col"] == 2


def test_splat_integer_column(oracle):
    # step==1 integer column: lo = binAt(x+1)-1, hi = binAt(n) → split halfway between the integers
    d = np.array([[desc(0.0, 4.0, 1.0, 4, is_int=1)]], abi.COL_DESC)
    h = hist_from([10, 10, 10, 10], [0, 0, 10, 10], [0, 0, 10, 10])
    assert oracle.find_splits(h, d, 1.0, 0.0)[0]["splat"] == np.float32(1.5)
    # wide integer column (step != 1): floor/ceil around the bin edge, the H2O "48.5" example
    d = np.array([[desc(0.0, 8.0, 0.5, 4, is_int=1)]], abi.COL_DESC)
    d[0, 0]["min"] = 0.4
    s = oracle.find_splits(h, d, 1.0, 0.0)[0]
    assert s["splat"] == np.float32((math.floor(4.4) + math.ceil(4.4)) / 2)


# ---------------------------------------------------------------- a8 / a6
def test_next_level_and_routing(oracle):
    p = abi.Params.make(nbins=4, min_rows=1.0)
    descs = np.array([[desc(0.0, 8.0, 0.5, 4), desc(-1.0, 1.0, 2.0, 4)]], abi.COL_DESC)
    h = np.concatenate([hist_from([10, 10, 10, 10], [0, 0, 10, 10], [0, 0, 10, 10])] * 2, axis=1)
    h[0, 1, :4, 1] = [5, 5, 5, 5.5]
    h[0, 1, :4, 2] = [5, 5, 5, 6]
    splits = oracle.find_splits(h, descs, 1.0, 0.0)
    assert splits[0]["col"] == 0 and splits[0]["splat"] == 4.0
    # perfect children (se == 0) get no histograms at all → both sides become leaves
    child, ids, n = oracle.next_level(splits, descs, np.array([[0.25, -0.5]], np.float32), np.array([[7.5, 0.75]], np.float32), p)
    assert n == 0 and ids.tolist() == [-1, -1]
    splits[0]["se_left"] = 1.0
    splits[0]["se_right"] = 1.0
    child, ids, n = oracle.next_level(splits, descs, np.array([[0.25, -0.5]], np.float32), np.array([[7.5, 0.75]], np.float32), p)
    assert n == 2 and ids.tolist() == [0, 1]
    # split column: left [min2, splat), right [splat, maxIn+ulp); other column: parent's observed range
    assert child[0, 0]["min"] == 0.25 and child[0, 0]["maxex"] == 4.0
    assert child[1, 0]["min"] == 4.0 and child[1, 0]["maxex"] == np.nextafter(7.5, np.inf)
    for k in (0, 1):
        assert child[k, 1]["min"] == -0.5 and child[k, 1]["maxex"] == np.nextafter(0.75, np.inf)
        assert child[k, 1]["step"] == 4 / (child[k, 1]["maxex"] + 0.5)
    # a column observed constant in the parent is dropped
    child, ids, n = oracle.next_level(splits, descs, np.array([[0.25, 0.5]], np.float32), np.array([[7.5, 0.5]], np.float32), p)
    assert child[0, 1]["nbin"] == 0 and child[1, 1]["nbin"] == 0
    # min_rows gate per side
    p2 = abi.Params.make(nbins=4, min_rows=25.0)
    _, ids, n = oracle.next_level(splits, descs, np.array([[0.25, -0.5]], np.float32), np.array([[7.5, 0.75]], np.float32), p2)
    assert n == 0
    # routing: x >= splat → right; NaN follows na_dir; rows of unsplit nodes and of child-less sides retire
    X = np.array([[3.99, 4.0, np.nan, 1.0, 9.0, 0, 0, 0], [0] * 8], np.float32)
    nid = np.array([0, 0, 0, 1, -1], np.int32)
    leaf = np.full(5, -1, np.int32)
    sp = np.zeros(2, abi.SPLIT)
    sp[0] = splits[0]
    sp[0]["na_dir"] = abi.NA_LEFT
    sp[1]["col"] = -1
    oracle.apply_splits(X, 5, sp, np.array([5, -1, -1, -1], np.int32), 10, nid, leaf)
    assert nid.tolist() == [5, -1, 5, -1, -1]
    assert leaf.tolist() == [-1, 10 * 3 + 2, -1, 11 * 3, -1]


def test_train_tree_invariants(oracle):
    rows, cols = 20000, 28
    X, y, _ = oracle.synth(abi.SHAPE_HIGGS28, 20261106, 0, rows, cols)
    p = abi.Params.make(nbins=64, max_depth=5)
    levels, nid, leaf = oracle.train_tree(X, rows, y, None, p)
    assert len(levels) == 5
    assert [lv["descs"].shape[0] for lv in levels][:3] == [1, 2, 4]  # informative y: real splits at the top
    for lv in levels:
        n_nodes = lv["descs"].shape[0]
        cnt = np.bincount(lv["nid"][lv["nid"] >= 0], minlength=n_nodes)
        for g in range(n_nodes):
            for c in range(cols):
                if lv["descs"][g, c]["nbin"] > 0:
                    assert lv["hist"][g, c, :, 0].sum() == cnt[g]  # every active row lands in exactly one bin
            s = lv["splits"][g]
            if s["col"] >= 0:
                assert s["n_left"] + s["n_right"] == cnt[g]
                assert s["se_left"] + s["se_right"] < s["se_before"]
    assert (nid < 0).all() and (leaf >= 0).all()  # after the last level every row sits in a leaf


# ---------------------------------------------------------------- a9
def test_gbm_passes_paper_case(oracle):
    label = np.array([1.0, 2.0, 4.0, 0.0], np.float32)
    f = np.array([0.5, 0.5, 1.0, 0.0], np.float32)
    res = np.zeros(4, np.float32)
    oracle.lib.h2ok_oracle_gbm_residual(abi.DIST_GAUSSIAN, abi.ptr(label), abi.ptr(f), None, 4, 1.0, abi.ptr(res), None)
    assert res.tolist() == [0.5, 1.5, 3.0, 0.0]
    # bernoulli: mu(0) = 0.5 → residual label - 0.5, denominator w*mu*(1-mu) = 0.25 (x 2^28)
    den = np.zeros(4, np.int32)
    f0 = np.zeros(4, np.float32)
    lab = np.array([1, 0, 1, 1], np.float32)
    oracle.lib.h2ok_oracle_gbm_residual(abi.DIST_BERNOULLI, abi.ptr(lab), abi.ptr(f0), None, 4, 2.0**28, abi.ptr(res), abi.ptr(den))
    assert res.tolist() == [0.5, -0.5, 0.5, 0.5] and den.tolist() == [2**26] * 4
    # GammaPass: integer sums per leaf code; rows with a negative code are ignored
    leaf = np.array([3, 3, 7, -1], np.int32)
    numq = np.array([10, -4, 5, 99], np.int32)
    num = np.zeros(9, np.int64)
    dn = np.zeros(9, np.int64)
    oracle.lib.h2ok_oracle_gbm_gamma_pass(abi.ptr(leaf), abi.ptr(numq), None, 4, 9, abi.ptr(num), abi.ptr(dn))
    assert num.tolist() == [0, 0, 0, 6, 0, 0, 0, 5, 0] and dn.tolist() == [0, 0, 0, 2, 0, 0, 0, 1, 0]
    import oracle_lib

    gamma = oracle_lib.gbm_gamma(num, dn, 2.0, 1.0)
    assert gamma.tolist() == [0, 0, 0, 1.5, 0, 0, 0, 2.5, 0]
    ff = np.array([1.0, 1.0, 1.0, 1.0], np.float32)
    oracle.lib.h2ok_oracle_gbm_add_gamma(abi.ptr(leaf), 4, abi.ptr(gamma), 9, 0.5, abi.ptr(ff))
    assert ff.tolist() == [1.75, 1.75, 2.25, 1.0]


def test_gbm_multinomial_residuals_paper_case(oracle):
    """K-class residual pass: zero scores give p = 1/K exactly; a score of ln 2 against zeros gives (1/2, 1/4, 1/4);
    residuals of a row sum to zero and the GammaPass denominator is |res| (1 - |res|)."""
    K, rows = 3, 4
    label = np.array([0, 1, 2, 1], np.float32)
    f = np.zeros((K, rows), np.float32)
    f[0, 3] = np.float32(np.log(2.0))
    res = np.zeros((K, rows), np.float32)
    den = np.zeros((K, rows), np.int32)
    oracle.lib.h2ok_oracle_gbm_residual_multi(K, abi.ptr(label), abi.ptr(f), rows, None, rows, 2.0**28, abi.ptr(res), abi.ptr(den))
    third = np.float32(1.0) / np.float32(3.0)
    for r in range(3):
        for k in range(K):
            want = (np.float32(1.0) if label[r] == k else np.float32(0.0)) - third
            assert res[k, r] == want
            a = abs(float(want))
            assert den[k, r] == int(np.rint(np.float64(np.float32(abs(want))) * np.float64(np.float32(1.0) - np.float32(abs(want))) * 2.0**28))
    assert res[:, 3] == pytest.approx([-0.5, 0.75, -0.25], abs=1e-6)
    assert abs(float(res[:, 3].astype(np.float64).sum())) < 1e-6
    assert den[:, 3] == pytest.approx([0.25 * 2**28, 0.1875 * 2**28, 0.1875 * 2**28], rel=1e-5)


# ---------------------------------------------------------------- categorical columns (is_int == 2)
def _bits_set(sp):
    return {32 * w + b for w in range(32) for b in range(32) if (int(sp["cat_bits"][w]) >> b) & 1}


def test_categorical_split_paper_case(oracle):
    # six categories (codes 3..8) whose means alternate: no threshold separates them, the sorted scan does
    cnt = [10, 10, 10, 10, 10, 10]
    ys = [0, 10, 0, 10, 0, 10]
    h = hist_from(cnt, ys, ys)
    d = np.array([[desc(3.0, 9.0, 1.0, 6, is_int=2)]], abi.COL_DESC)
    s = oracle.find_splits(h, d, 10.0, 1e-5)[0]
    assert s["col"] == 0 and s["equal"] == 2 and s["bin"] == 3
    assert s["se_left"] == 0 and s["se_right"] == 0 and s["pred_left"] == 0 and s["pred_right"] == 1
    assert _bits_set(s) == {4, 6, 8}  # the categories with mean 1 go right
    # the same column as an ordinary integer column: an ordered split, far worse
    d1 = np.array([[desc(3.0, 9.0, 1.0, 6, is_int=1)]], abi.COL_DESC)
    s1 = oracle.find_splits(h, d1, 10.0, 1e-5)[0]
    assert s1["equal"] == 0 and s1["se_left"] + s1["se_right"] > 10
    # fewer than four bins: ordered scan (no bit set)
    h3 = hist_from(cnt[:3], ys[:3], ys[:3])
    d3 = np.array([[desc(3.0, 6.0, 1.0, 3, is_int=2)]], abi.COL_DESC)
    assert oracle.find_splits(h3, d3, 10.0, 1e-5)[0]["equal"] == 0
    # unseen categories (empty bins) follow the majority of the seen ones; ties go left
    h = hist_from([10, 0, 10, 10, 0, 10], [0, 0, 10, 10, 0, 10], [0, 0, 10, 10, 0, 10])
    s = oracle.find_splits(h, d, 10.0, 1e-5)[0]
    assert s["equal"] == 2 and _bits_set(s) == {5, 6, 8} | {4, 7}  # one seen category left, three right: empties right
    h = hist_from([10, 0, 10, 10, 0, 10], [0, 0, 0, 10, 0, 10], [0, 0, 0, 10, 0, 10])
    assert _bits_set(oracle.find_splits(h, d, 10.0, 1e-5)[0]) == {6, 8}  # two each: empties left


def test_categorical_sorted_scan_finds_the_best_partition(oracle):
    """The property the sort is there for: for squared error, the best of all 2^(k-1) - 1 two-set partitions of k
    categories is a cut of the categories sorted by mean.  Brute force over every partition, several random cases."""
    rng = np.random.default_rng(5)
    for case in range(20):
        k = int(rng.integers(4, 9))
        cnt = rng.integers(1, 20, k).astype(np.float64)
        mean = rng.integers(-8, 9, k) / 4.0
        ys = cnt * mean
        yys = cnt * mean * mean + rng.integers(0, 5, k) / 4.0
        h = hist_from(cnt, ys, yys)
        d = np.array([[desc(0.0, float(k), 1.0, k, is_int=2)]], abi.COL_DESC)
        s = oracle.find_splits(h, d, 1.0, 0.0)[0]

        def se(mask):
            w, wy, wyy = cnt[mask].sum(), ys[mask].sum(), yys[mask].sum()
            return wyy - wy * wy / w

        best = min(se(np.array([(m >> i) & 1 == 1 for i in range(k)])) + se(np.array([(m >> i) & 1 == 0 for i in range(k)]))
                   for m in range(1, 2 ** k - 1))
        if s["col"] < 0:
            continue
        assert s["equal"] == 2
        assert abs((s["se_left"] + s["se_right"]) - best) <= 1e-9 * max(1.0, abs(best))
        right = np.array([i in _bits_set(s) for i in range(k)])
        assert abs(se(right) - s["se_right"]) <= 1e-9 * max(1.0, s["se_right"]) and cnt[right].sum() == s["n_right"]


def test_categorical_next_level_and_routing(oracle):
    p = abi.Params.make(nbins=8, min_rows=1.0)
    descs = np.array([[desc(0.0, 6.0, 1.0, 6, is_int=2)]], abi.COL_DESC)
    sp = np.zeros(1, abi.SPLIT)
    sp[0]["col"], sp[0]["equal"], sp[0]["na_dir"] = 0, 2, abi.NA_LEFT
    sp[0]["n_left"] = sp[0]["n_right"] = 10
    sp[0]["se_left"] = sp[0]["se_right"] = 1.0
    sp[0]["cat_bits"][0] = (1 << 1) | (1 << 4)
    # both children keep the parent's observed range of the split column: nothing narrower is known
    child, ids, n = oracle.next_level(sp, descs, np.array([[0.0]], np.float32), np.array([[5.0]], np.float32), p)
    assert n == 2
    for k in (0, 1):
        assert child[k, 0]["min"] == 0.0 and child[k, 0]["maxex"] == 6.0 and child[k, 0]["nbin"] == 6 and child[k, 0]["is_int"] == 2
    X = np.array([[0, 1, 2, 4, np.nan, 2000.0, -3.0, 5]], np.float32)
    nid = np.zeros(8, np.int32)
    leaf = np.full(8, -1, np.int32)
    oracle.apply_splits(X, 8, sp, np.array([0, 1], np.int32), 0, nid, leaf)
    assert nid.tolist() == [0, 1, 0, 1, 0, 0, 0, 0]  # NaN and codes outside 0..1023 follow na_dir (left)
