"""ctypes binding of oracle/liboracle.so — TEST INFRASTRUCTURE ONLY (tests/, smoke(), bench cpu_baseline).

Builds the library with `make -C oracle` if it is missing (gcc only, no GPU needed).
"""
import ctypes as C
import os
import subprocess

import numpy as np

from h2ok_hist import abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = None

i32, i64, u64, f64 = C.c_int32, C.c_int64, C.c_uint64, C.c_double
vp = C.c_void_p


def _build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])


class Oracle:
    def __init__(self, lib):
        self.lib = lib
        L = lib
        L.h2ok_oracle_splitmix64.restype = u64
        L.h2ok_oracle_splitmix64.argtypes = [u64]
        L.h2ok_oracle_gen_x.restype = C.c_float
        L.h2ok_oracle_gen_x.argtypes = [i32, u64, i64, i32, i32]
        L.h2ok_oracle_synth_fill.restype = None
        L.h2ok_oracle_synth_fill.argtypes = [i32, u64, i64, i64, i32, i64, i32, i32, vp, vp, vp]
        L.h2ok_oracle_make_desc.restype = i32
        L.h2ok_oracle_make_desc.argtypes = [f64, f64, i32, i32, vp]
        L.h2ok_oracle_find_maxex.restype = f64
        L.h2ok_oracle_find_maxex.argtypes = [f64, i32]
        L.h2ok_oracle_col_ranges.restype = None
        L.h2ok_oracle_col_ranges.argtypes = [vp, i64, i64, i32, vp, vp]
        L.h2ok_oracle_root_descs.restype = i32
        L.h2ok_oracle_root_descs.argtypes = [vp, i64, i64, i32, C.POINTER(abi.Params), vp, vp]
        L.h2ok_oracle_bin.restype = i32
        L.h2ok_oracle_bin.argtypes = [C.c_float, vp]
        L.h2ok_oracle_bin_many.restype = None
        L.h2ok_oracle_bin_many.argtypes = [vp, i64, vp, vp]
        L.h2ok_oracle_target_absmax.restype = None
        L.h2ok_oracle_target_absmax.argtypes = [vp, vp, i64, vp]
        L.h2ok_oracle_choose_quant.restype = None
        L.h2ok_oracle_choose_quant.argtypes = [vp, i32, i32, C.POINTER(abi.Quant)]
        L.h2ok_oracle_quantize_rows.restype = None
        L.h2ok_oracle_quantize_rows.argtypes = [vp, vp, i64, C.POINTER(abi.Quant), vp, vp, vp]
        L.h2ok_oracle_build_level.restype = None
        L.h2ok_oracle_build_level.argtypes = [vp, i64, i64, i32, vp, vp, vp, i32, vp, i32, vp, vp, vp, vp, i32]
        L.h2ok_oracle_build_level_q.restype = None
        L.h2ok_oracle_build_level_q.argtypes = [vp, i64, i64, i32, vp, vp, vp, vp, i32, vp, i32, vp, vp, vp]
        L.h2ok_oracle_finalize.restype = None
        L.h2ok_oracle_finalize.argtypes = [vp, i64, i32, C.POINTER(abi.Quant), vp]
        L.h2ok_oracle_find_splits.restype = None
        L.h2ok_oracle_find_splits.argtypes = [vp, vp, i32, i32, i32, f64, f64, vp]
        L.h2ok_oracle_next_level.restype = None
        L.h2ok_oracle_next_level.argtypes = [vp, vp, vp, vp, i32, i32, C.POINTER(abi.Params), i32, vp, vp, vp]
        L.h2ok_oracle_apply_splits.restype = None
        L.h2ok_oracle_apply_splits.argtypes = [vp, i64, i64, vp, vp, i32, i32, vp, vp]
        L.h2ok_oracle_gbm_residual.restype = None
        L.h2ok_oracle_gbm_residual.argtypes = [i32, vp, vp, vp, i64, f64, vp, vp]
        L.h2ok_oracle_gbm_residual_multi.restype = None
        L.h2ok_oracle_gbm_residual_multi.argtypes = [i32, vp, vp, i64, vp, i64, f64, vp, vp]
        L.h2ok_oracle_gbm_gamma_pass.restype = None
        L.h2ok_oracle_gbm_gamma_pass.argtypes = [vp, vp, vp, i64, i32, vp, vp]
        L.h2ok_oracle_gbm_add_gamma.restype = None
        L.h2ok_oracle_gbm_add_gamma.argtypes = [vp, i64, vp, i32, f64, vp]
        L.h2ok_oracle_max_threads.restype = i32

    # ---- generator ----
    def synth(self, shape, seed, row0, rows, cols, na_permille=0, weight_mode=0, ld=None):
        ld = ld or ((rows + 3) // 4 * 4)
        X = np.zeros((cols, ld), np.float32)
        y = np.zeros(rows, np.float32)
        w = np.zeros(rows, np.float32) if weight_mode else None
        self.lib.h2ok_oracle_synth_fill(shape, seed, row0, rows, cols, ld, na_permille, weight_mode,
                                        abi.ptr(X), abi.ptr(y), abi.ptr(w))
        return X, y, w

    # ---- descriptors ----
    def make_desc(self, mn, maxex, is_int, nbins):
        d = np.zeros(1, abi.COL_DESC)
        ok = self.lib.h2ok_oracle_make_desc(mn, maxex, is_int, nbins, abi.ptr(d))
        return d[0], ok

    def find_maxex(self, maxin, is_int):
        return self.lib.h2ok_oracle_find_maxex(maxin, is_int)

    def col_ranges(self, X, rows):
        cols, ld = X.shape
        mn = np.zeros(cols, np.float32)
        mx = np.zeros(cols, np.float32)
        self.lib.h2ok_oracle_col_ranges(abi.ptr(X), ld, rows, cols, abi.ptr(mn), abi.ptr(mx))
        return mn, mx

    def root_descs(self, X, rows, params, is_int=None):
        cols, ld = X.shape
        out = np.zeros(cols, abi.COL_DESC)
        ii = None if is_int is None else np.ascontiguousarray(is_int, np.int32)
        rc = self.lib.h2ok_oracle_root_descs(abi.ptr(X), ld, rows, cols, C.byref(params), abi.ptr(ii), abi.ptr(out))
        assert rc == 0
        return out

    def bin(self, x, desc):
        d = np.array([desc], abi.COL_DESC)
        return self.lib.h2ok_oracle_bin(C.c_float(x), abi.ptr(d))

    def bin_many(self, x, desc):
        x = np.ascontiguousarray(x, np.float32)
        d = np.array([desc], abi.COL_DESC)
        out = np.zeros(x.size, np.int32)
        self.lib.h2ok_oracle_bin_many(abi.ptr(x), x.size, abi.ptr(d), abi.ptr(out))
        return out

    # ---- quantisation ----
    def choose_quant(self, y, w, qbits=25):
        m = np.zeros(3, np.float64)
        self.lib.h2ok_oracle_target_absmax(abi.ptr(y), abi.ptr(w), y.size, abi.ptr(m))
        q = abi.Quant()
        self.lib.h2ok_oracle_choose_quant(abi.ptr(m), 1 if w is not None else 0, qbits, C.byref(q))
        return q, m

    def quantize_rows(self, y, w, q):
        n = y.size
        wq = np.zeros(n, np.int32) if q.weighted else None
        wyq = np.zeros(n, np.int32)
        wyyq = np.zeros(n, np.int32)
        self.lib.h2ok_oracle_quantize_rows(abi.ptr(y), abi.ptr(w), n, C.byref(q), abi.ptr(wq), abi.ptr(wyq), abi.ptr(wyyq))
        return wq, wyq, wyyq

    # ---- level ----
    def build_level(self, X, rows, y, w, nid, descs, max_bins, threads=1, want_abs=True):
        cols, ld = X.shape
        descs = np.ascontiguousarray(descs, abi.COL_DESC).reshape(-1, cols)
        n_nodes = descs.shape[0]
        hist = np.zeros((n_nodes, cols, max_bins + 1, 3), np.float64)
        absy = np.zeros((n_nodes, cols, max_bins + 1), np.float64) if want_abs else None
        min2 = np.zeros((n_nodes, cols), np.float32)
        maxin = np.zeros((n_nodes, cols), np.float32)
        self.lib.h2ok_oracle_build_level(abi.ptr(X), ld, rows, cols, abi.ptr(y), abi.ptr(w), abi.ptr(nid), n_nodes,
                                         abi.ptr(descs), max_bins, abi.ptr(hist), abi.ptr(absy), abi.ptr(min2),
                                         abi.ptr(maxin), threads)
        return hist, absy, min2, maxin

    def build_level_q(self, X, rows, nid, wq, wyq, wyyq, descs, max_bins):
        cols, ld = X.shape
        descs = np.ascontiguousarray(descs, abi.COL_DESC).reshape(-1, cols)
        n_nodes = descs.shape[0]
        acc = np.zeros((n_nodes, cols, max_bins + 1, 3), np.int64)
        min2 = np.zeros((n_nodes, cols), np.float32)
        maxin = np.zeros((n_nodes, cols), np.float32)
        self.lib.h2ok_oracle_build_level_q(abi.ptr(X), ld, rows, cols, abi.ptr(nid), abi.ptr(wq), abi.ptr(wyq),
                                           abi.ptr(wyyq), n_nodes, abi.ptr(descs), max_bins, abi.ptr(acc),
                                           abi.ptr(min2), abi.ptr(maxin))
        return acc, min2, maxin

    def finalize(self, acc, q, max_bins):
        hist = np.zeros(acc.shape, np.float64)
        self.lib.h2ok_oracle_finalize(abi.ptr(acc), acc.size, max_bins, C.byref(q), abi.ptr(hist))
        return hist

    def find_splits(self, hist, descs, min_rows=10.0, msi=1e-5):
        n_nodes, cols, nb1, _ = hist.shape
        hist = np.ascontiguousarray(hist, np.float64)
        descs = np.ascontiguousarray(descs, abi.COL_DESC).reshape(n_nodes, cols)
        out = np.zeros(n_nodes, abi.SPLIT)
        self.lib.h2ok_oracle_find_splits(abi.ptr(hist), abi.ptr(descs), n_nodes, cols, nb1 - 1, min_rows, msi, abi.ptr(out))
        return out

    def next_level(self, splits, descs, min2, maxin, params, depth=0):
        descs = np.ascontiguousarray(descs, abi.COL_DESC)
        n_nodes, cols = descs.shape
        splits = np.ascontiguousarray(splits, abi.SPLIT)
        child = np.zeros((2 * n_nodes, cols), abi.COL_DESC)
        ids = np.zeros(2 * n_nodes, np.int32)
        n = C.c_int32(0)
        min2 = np.ascontiguousarray(min2, np.float32)
        maxin = np.ascontiguousarray(maxin, np.float32)
        self.lib.h2ok_oracle_next_level(abi.ptr(splits), abi.ptr(descs), abi.ptr(min2), abi.ptr(maxin), n_nodes, cols,
                                        C.byref(params), depth, abi.ptr(child), abi.ptr(ids), C.byref(n))
        return child[: n.value].copy(), ids, n.value

    def apply_splits(self, X, rows, splits, child_ids, gid_base, nid, leaf=None):
        cols, ld = X.shape
        splits = np.ascontiguousarray(splits, abi.SPLIT)
        child_ids = np.ascontiguousarray(child_ids, np.int32)
        self.lib.h2ok_oracle_apply_splits(abi.ptr(X), ld, rows, abi.ptr(splits), abi.ptr(child_ids), gid_base,
                                          splits.size, abi.ptr(nid), abi.ptr(leaf))

    def max_threads(self):
        return self.lib.h2ok_oracle_max_threads()

    # ---- whole tree on the CPU (used as the checker for the level loop) ----
    def train_tree(self, X, rows, y, w, params, is_int=None, threads=1, on_level=None):
        cols = X.shape[0]
        descs = self.root_descs(X, rows, params, is_int).reshape(1, cols)
        nid = np.zeros(rows, np.int32)
        leaf = np.full(rows, -1, np.int32)
        gid_base = 0
        levels = []
        max_bins = max(params.nbins, params.nbins_top_level)
        for depth in range(params.max_depth):
            n_nodes = descs.shape[0]
            hist, absy, min2, maxin = self.build_level(X, rows, y, w, nid, descs, max_bins, threads)
            splits = self.find_splits(hist, descs, params.min_rows, params.min_split_improvement)
            if depth + 1 < params.max_depth:
                child, ids, nch = self.next_level(splits, descs, min2, maxin, params, depth)
            else:
                child, ids, nch = np.zeros((0, cols), abi.COL_DESC), np.full(2 * n_nodes, -1, np.int32), 0
            levels.append(dict(descs=descs, hist=hist, abs_wy=absy, min2=min2, maxin=maxin, splits=splits,
                               child_ids=ids, nid=nid.copy(), gid_base=gid_base))
            if on_level:
                on_level(depth, levels[-1])
            self.apply_splits(X, rows, splits, ids, gid_base, nid, leaf)
            gid_base += n_nodes
            descs = child
            if nch == 0:
                break
        return levels, nid, leaf


def train_tree_q(oracle, X, rows, y, w, params, is_int=None, qbits=25, on_level=None):
    """The level loop over the FIXED-POINT targets (every channel an exact integer): what the CUDA path must equal
    bit for bit.  Returns (levels, nid, leaf, quant); a level holds descs, acc, hist, min2, maxin, splits, child_ids,
    child_descs, nid (before routing), gid_base."""
    cols = X.shape[0]
    descs = oracle.root_descs(X, rows, params, is_int).reshape(1, cols)
    q, _ = oracle.choose_quant(y[:rows], None if w is None else w[:rows], qbits)
    wq, wyq, wyyq = oracle.quantize_rows(y[:rows], None if w is None else w[:rows], q)
    nid = np.zeros(rows, np.int32)
    leaf = np.full(rows, -1, np.int32)
    gid_base, levels = 0, []
    max_bins = max(params.nbins, params.nbins_top_level)
    for depth in range(params.max_depth):
        n = descs.shape[0]
        acc, min2, maxin = oracle.build_level_q(X, rows, nid, wq, wyq, wyyq, descs, max_bins)
        hist = oracle.finalize(acc, q, max_bins)
        splits = oracle.find_splits(hist, descs, params.min_rows, params.min_split_improvement)
        if depth + 1 < params.max_depth:
            child, ids, nch = oracle.next_level(splits, descs, min2, maxin, params, depth)
        else:
            child, ids, nch = np.zeros((0, cols), abi.COL_DESC), np.full(2 * n, -1, np.int32), 0
        levels.append(dict(descs=descs, acc=acc, hist=hist, min2=min2, maxin=maxin, splits=splits, child_ids=ids,
                           child_descs=child, nid=nid.copy(), gid_base=gid_base, depth=depth))
        if on_level:
            on_level(depth, levels[-1])
        oracle.apply_splits(X, rows, splits, ids, gid_base, nid, leaf)
        gid_base += n
        descs = child
        if nch == 0:
            break
    return levels, nid, leaf, q


# ---- host restatement of h2ok_hist_loop_checksum (csrc/tree_driver.cu) --------------------------------------------
_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)
SALT_ACC, SALT_SPLIT, SALT_LEAF = 0xACC0, 0x5911, 0x1EAF


def splitmix64_np(z):
    z = np.asarray(z, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = z + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def _hash_sum(values_i64, idx0, salt):
    v = np.ascontiguousarray(values_i64, dtype=np.int64).reshape(-1).view(np.uint64)
    total = np.uint64(0)
    step = 1 << 22
    with np.errstate(over="ignore"):
        for a in range(0, v.size, step):
            i = np.arange(a, min(a + step, v.size), dtype=np.uint64) + np.uint64(idx0)
            total = total + np.sum(splitmix64_np(i ^ np.uint64(salt)) * v[a:a + step], dtype=np.uint64)
    return int(total)


def tree_checksum(levels, leaf, row0=0):
    """[acc, splits, leaf] sums of a tree given its levels (each with acc, splits, gid_base, depth) and leaf codes."""
    acc_sum = split_sum = 0
    for lv in levels:
        salt = int(splitmix64_np(np.uint64(SALT_ACC + lv["depth"])))
        acc_sum = (acc_sum + _hash_sum(lv["acc"], 0, salt)) & 0xFFFFFFFFFFFFFFFF
        sp = lv["splits"]
        f = np.stack([sp["col"].astype(np.int64), sp["bin"].astype(np.int64), sp["na_dir"].astype(np.int64),
                      sp["splat"].view(np.uint32).astype(np.int64)], axis=1)
        split_sum = (split_sum + _hash_sum(f, 4 * lv["gid_base"], SALT_SPLIT)) & 0xFFFFFFFFFFFFFFFF
        if "equal" in sp.dtype.names:  # categorical splits: their bit sets, word by word
            with np.errstate(over="ignore"):
                for g in np.nonzero((sp["col"] >= 0) & (sp["equal"] == 2))[0]:
                    b = np.uint64(4 * (lv["gid_base"] + int(g)) + 1)
                    w = np.arange(1, 33, dtype=np.uint64) << np.uint64(32)
                    h = splitmix64_np(b ^ np.uint64(SALT_SPLIT) ^ w) * sp["cat_bits"][g].astype(np.uint64)
                    split_sum = (split_sum + int(np.sum(h, dtype=np.uint64))) & 0xFFFFFFFFFFFFFFFF
    return [acc_sum, split_sum, _hash_sum(np.asarray(leaf, np.int32).astype(np.int64), row0, SALT_LEAF)]


def gbm_gamma(num, den, scale_num, scale_den):
    """Leaf values from the integer GammaPass sums (same expression as h2ok_hist.gbm.leaf_gamma)."""
    num = num.astype(np.float64) / scale_num
    den = den.astype(np.float64) / scale_den
    return np.where(den != 0, num / np.where(den != 0, den, 1.0), 0.0)


class OracleGBM:
    """The a9 loop on the CPU with the fixed-point level walk — the bit-exact checker for h2ok_hist.gbm.GBMTrainer."""

    def __init__(self, oracle, X, rows, label, w, params, dist, learn_rate, qbits=25, is_int=None, n_classes=1):
        self.o, self.X, self.rows, self.label, self.w = oracle, X, rows, label, w
        self.params, self.dist, self.lr, self.qbits, self.is_int = params, dist, learn_rate, qbits, is_int
        self.cols = X.shape[0]
        self.root = oracle.root_descs(X, rows, params, is_int)
        self.K = n_classes
        self.f = np.zeros(rows, np.float32) if dist != abi.DIST_MULTINOMIAL else np.zeros((n_classes, rows), np.float32)
        self.trees = []

    def init_f(self, f0):
        self.f[:] = np.float32(f0)

    def tree_q(self, res):
        """One tree on targets `res` with the quantised walk; returns leaf codes, quant, wyq, total nodes."""
        o, X, rows, p = self.o, self.X, self.rows, self.params
        q, _ = o.choose_quant(res, self.w, self.qbits)
        wq, wyq, wyyq = o.quantize_rows(res, self.w, q)
        descs = self.root.reshape(1, self.cols).copy()
        nid = np.zeros(rows, np.int32)
        leaf = np.full(rows, -1, np.int32)
        gid, nb = 0, max(p.nbins, p.nbins_top_level)
        levels = []
        for depth in range(p.max_depth):
            n = descs.shape[0]
            acc, min2, maxin = o.build_level_q(X, rows, nid, wq, wyq, wyyq, descs, nb)
            splits = o.find_splits(o.finalize(acc, q, nb), descs, p.min_rows, p.min_split_improvement)
            if depth + 1 < p.max_depth:
                child, ids, nch = o.next_level(splits, descs, min2, maxin, p, depth)
            else:
                child, ids, nch = None, np.full(2 * n, -1, np.int32), 0
            o.apply_splits(X, rows, splits, ids, gid, nid, leaf)
            levels.append((splits, ids))
            gid += n
            descs = child
            if nch == 0:
                break
        return leaf, q, wq, wyq, gid, levels

    def step_multinomial(self):
        """K trees per iteration on the softmax residuals (GBM.java buildNextKTrees, multinomial)."""
        o, rows, K = self.o, self.rows, self.K
        den_scale = 2.0**28
        res = np.zeros((K, rows), np.float32)
        denq = np.zeros((K, rows), np.int32)
        o.lib.h2ok_oracle_gbm_residual_multi(K, abi.ptr(self.label), abi.ptr(self.f), rows, abi.ptr(self.w), rows, den_scale,
                                             abi.ptr(res), abi.ptr(denq))
        out = []
        for k in range(K):
            leaf, q, wq, wyq, n_tree_nodes, levels = self.tree_q(np.ascontiguousarray(res[k]))
            n_codes = 3 * n_tree_nodes
            num = np.zeros(n_codes, np.int64)
            den = np.zeros(n_codes, np.int64)
            dk = np.ascontiguousarray(denq[k])
            o.lib.h2ok_oracle_gbm_gamma_pass(abi.ptr(leaf), abi.ptr(wyq), abi.ptr(dk), rows, n_codes, abi.ptr(num), abi.ptr(den))
            gamma = (K - 1.0) / K * gbm_gamma(num, den, q.scale_wy, den_scale)
            fk = np.ascontiguousarray(self.f[k])
            o.lib.h2ok_oracle_gbm_add_gamma(abi.ptr(leaf), rows, abi.ptr(gamma), n_codes, self.lr, abi.ptr(fk))
            self.f[k] = fk
            self.trees.append(dict(levels=levels, gamma=gamma, leaf=leaf, cls=k))
            out.append(gamma)
        return out

    def step(self):
        if self.dist == abi.DIST_MULTINOMIAL:
            return self.step_multinomial()
        o, rows = self.o, self.rows
        den_scale = 2.0**28 if self.dist == abi.DIST_BERNOULLI else None
        res = np.zeros(rows, np.float32)
        # gaussian: the denominator is the weight (quantised with the histogram's own scale) or the row count
        denq = None
        if self.dist == abi.DIST_BERNOULLI:
            denq = np.zeros(rows, np.int32)
            o.lib.h2ok_oracle_gbm_residual(self.dist, abi.ptr(self.label), abi.ptr(self.f), abi.ptr(self.w), rows, den_scale,
                                           abi.ptr(res), abi.ptr(denq))
        else:
            o.lib.h2ok_oracle_gbm_residual(self.dist, abi.ptr(self.label), abi.ptr(self.f), abi.ptr(self.w), rows, 1.0,
                                           abi.ptr(res), None)
        leaf, q, wq, wyq, n_tree_nodes, levels = self.tree_q(res)
        if self.dist == abi.DIST_GAUSSIAN:
            denq, den_scale = (wq, q.scale_w) if self.w is not None else (None, 1.0)
        n_codes = 3 * n_tree_nodes
        num = np.zeros(n_codes, np.int64)
        den = np.zeros(n_codes, np.int64)
        o.lib.h2ok_oracle_gbm_gamma_pass(abi.ptr(leaf), abi.ptr(wyq), abi.ptr(denq), rows, n_codes, abi.ptr(num), abi.ptr(den))
        gamma = gbm_gamma(num, den, q.scale_wy, den_scale)
        o.lib.h2ok_oracle_gbm_add_gamma(abi.ptr(leaf), rows, abi.ptr(gamma), n_codes, self.lr, abi.ptr(self.f))
        self.trees.append(dict(levels=levels, gamma=gamma, leaf=leaf))
        return gamma


def load():
    global _LIB
    if _LIB is None:
        path = os.path.join(ROOT, "oracle", "liboracle.so")
        src = os.path.join(ROOT, "oracle", "hist_oracle.c")
        if not os.path.exists(path) or (os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(path)):
            _build()
        _LIB = Oracle(C.CDLL(path))
    return _LIB
