"""Host-side mirror of the histogram path: one rank's resident shard (CudaShard) and the per-level loop
(LevelDriver) that H2O-3 runs as SharedTree.buildLayer → ScoreBuildHistogram2 (map over rows) → MRTask.reduce
(DHistogram.add) → DTree.findBestSplitPoint → Split.nextLevelHistos → score_decide.

PyTorch is plumbing only: device memory, the CUDA stream, and torch.distributed for the one real exchange step
(the histogram reduce).  All arithmetic happens in libh2ok_hist.so through the C ABI; there is no CPU path here.

Rows are sharded across ranks; descriptors, splits and child ids are replicated.  Because the
accumulators are integers, the all-reduce is exact and every rank derives bit-identical splits.
"""
import ctypes as C

import numpy as np
import torch
import torch.distributed as dist

from . import abi, capi

_I32_SIGN = -(2**31)
_I32_MAX = 2**31 - 1


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


class CudaShard:
    """One GPU's rows: column-major fp32 X [cols, ld], targets, row→node map, and the level buffers."""

    def __init__(self, device=0, max_level_nodes=128):
        self.lib = capi.load()
        self.device = torch.device("cuda", device)
        torch.cuda.set_device(self.device)
        h = C.c_void_p()
        self.lib.call("h2ok_hist_create", device, C.byref(h))
        self.ctx = h
        self.use_current_stream()
        self.max_level_nodes = max_level_nodes
        self.rows = self.cols = self.ld = 0
        self.X = self.y = self.w = None
        self.weighted = False
        self.quant = None
        self._level_cap = None
        self.launches = 0  # kernels of libh2ok_hist.so launched through this shard

    def close(self):
        if self.ctx is not None:
            self.lib.call("h2ok_hist_destroy", self.ctx)
            self.ctx = None

    def use_current_stream(self):
        """Kernels go on torch's current stream so torch.cuda.Event timing and NCCL ordering see them."""
        self.lib.call("h2ok_hist_set_stream", self.ctx, C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream), 0)

    # ---- data -------------------------------------------------------------------------------------------
    def _alloc_rows(self, rows, cols, weighted):
        self.rows, self.cols = rows, cols
        self.ld = (rows + 63) // 64 * 64
        dev = self.device
        self.X = torch.empty((cols, self.ld), dtype=torch.float32, device=dev)
        self.y = torch.empty(self.ld, dtype=torch.float32, device=dev)
        self.w = torch.empty(self.ld, dtype=torch.float32, device=dev) if weighted else None
        self.weighted = weighted
        self.nid = torch.zeros(self.ld, dtype=torch.int32, device=dev)
        self.leaf = torch.full((self.ld,), -1, dtype=torch.int32, device=dev)
        self.wq = torch.empty(self.ld, dtype=torch.int32, device=dev) if weighted else None
        self.wyq = torch.empty(self.ld, dtype=torch.int32, device=dev)
        self.wyyq = torch.empty(self.ld, dtype=torch.int32, device=dev)

    def set_data_synth(self, shape, seed, row0, rows, cols, na_permille=0, weight_mode=0):
        """Counter-based generator: this rank fills rows [row0,row0+rows) of the global matrix."""
        self._alloc_rows(rows, cols, weight_mode != 0)
        self.lib.call("h2ok_hist_synth_fill_dev", self.ctx, shape, seed, row0, rows, cols, self.ld, na_permille,
                      weight_mode, _ptr(self.X), _ptr(self.y), _ptr(self.w))

    def set_data_tensors(self, X, y, w=None):
        """X: [cols, rows] fp32 (host or device), copied into the padded column-major layout."""
        cols, rows = X.shape
        self._alloc_rows(rows, cols, w is not None)
        self.X[:, :rows].copy_(X, non_blocking=True)
        self.X[:, rows:].zero_()
        self.set_targets(y, w)

    def set_targets(self, y, w=None):
        self.y[: self.rows].copy_(y, non_blocking=True)
        if self.weighted:
            self.w[: self.rows].copy_(w, non_blocking=True)

    def synth_targets(self, seed, row0):
        self.lib.call("h2ok_hist_synth_fill_dev", self.ctx, abi.SHAPE_HIGGS28, seed, row0, self.rows, self.cols, self.ld,
                      0, 1 if self.weighted else 0, None, _ptr(self.y), _ptr(self.w))

    # ---- per-tree set-up ---------------------------------------------------------------------------------
    def col_ranges(self):
        mn = torch.empty(self.cols, dtype=torch.int32, device=self.device)
        mx = torch.empty(self.cols, dtype=torch.int32, device=self.device)
        self.lib.call("h2ok_hist_col_ranges_dev", self.ctx, _ptr(self.X), self.ld, self.rows, self.cols, _ptr(mn), _ptr(mx))
        return mn, mx

    def target_absmax(self):
        out = torch.empty(3, dtype=torch.float64, device=self.device)
        self.lib.call("h2ok_hist_target_absmax_dev", self.ctx, _ptr(self.y), _ptr(self.w), self.rows, _ptr(out))
        return out

    def quantize(self, quant):
        self.quant = quant
        self.lib.call("h2ok_hist_quantize_rows_dev", self.ctx, _ptr(self.y), _ptr(self.w), self.rows, C.byref(quant),
                      _ptr(self.wq), _ptr(self.wyq), _ptr(self.wyyq))

    def reset_nodes(self):
        self.nid.zero_()
        self.leaf.fill_(-1)

    # ---- level buffers -----------------------------------------------------------------------------------
    def ensure_level_buffers(self, max_nodes, max_bins):
        key = (max_nodes, max_bins, self.cols)
        if self._level_cap == key:
            return
        dev, cols, nb1 = self.device, self.cols, max_bins + 1
        self.descs = torch.zeros((max_nodes, cols, abi.COL_DESC.itemsize), dtype=torch.uint8, device=dev)
        self.child_descs = torch.zeros((2 * max_nodes, cols, abi.COL_DESC.itemsize), dtype=torch.uint8, device=dev)
        self.acc = torch.zeros((max_nodes, cols, nb1, 3), dtype=torch.int64, device=dev)
        self.hist = torch.zeros((max_nodes, cols, nb1, 3), dtype=torch.float64, device=dev)
        self.mn = torch.zeros((max_nodes, cols), dtype=torch.int32, device=dev)
        self.mx = torch.zeros((max_nodes, cols), dtype=torch.int32, device=dev)
        self.min2 = torch.zeros((max_nodes, cols), dtype=torch.float32, device=dev)
        self.maxin = torch.zeros((max_nodes, cols), dtype=torch.float32, device=dev)
        self.splits = torch.zeros((max_nodes, abi.SPLIT.itemsize), dtype=torch.uint8, device=dev)
        self.child_ids = torch.zeros(2 * max_nodes, dtype=torch.int32, device=dev)
        self.n_children = torch.zeros(1, dtype=torch.int32, device=dev)
        self._level_cap = key
        self.max_bins = max_bins

    def upload_descs(self, descs_np):
        d = np.ascontiguousarray(descs_np, abi.COL_DESC).reshape(-1, self.cols)
        n = d.shape[0]
        self.descs[:n].copy_(torch.from_numpy(d.view(np.uint8).reshape(n, self.cols, -1)))
        return n

    def adopt_children(self, n):
        """The children computed by next_level become the next level's descriptors (device-to-device)."""
        self.descs[:n].copy_(self.child_descs[:n])

    # ---- a3/a4: local histogram --------------------------------------------------------------------------
    def build_level(self, n_nodes, level_bins=0):
        self.lib.call("h2ok_hist_build_level_dev", self.ctx, _ptr(self.X), self.ld, self.rows, self.cols, _ptr(self.nid),
                      _ptr(self.wq), _ptr(self.wyq), _ptr(self.wyyq), n_nodes, _ptr(self.descs), self.max_bins, 1,
                      _ptr(self.acc), _ptr(self.mn), _ptr(self.mx))
        # init_enc + pack_rows + one hist_build per node group (the library's own plan)
        self.launches += 2 + self.lib.value("h2ok_hist_build_level_passes", n_nodes, self.cols, self.rows, self.max_bins)
        return self.acc[:n_nodes], self.mn[:n_nodes], self.mx[:n_nodes]

    def last_hist_ms(self):
        """Device time of the last build_level's hist_build launches (CUDA events inside the library)."""
        return float(self.stats().build_ms)

    def finalize_level(self, n_nodes, hist=True, ranges=True):
        """Integer accumulators -> fp64 histogram (hist) and encoded -> float observed ranges (ranges)."""
        self.lib.call("h2ok_hist_finalize_level_dev", self.ctx, _ptr(self.acc) if hist else None,
                      _ptr(self.mn) if ranges else None, _ptr(self.mx) if ranges else None, n_nodes, self.cols,
                      self.max_bins, C.byref(self.quant), _ptr(self.hist) if hist else None,
                      _ptr(self.min2) if ranges else None, _ptr(self.maxin) if ranges else None)
        self.launches += int(hist) + int(ranges)

    # ---- a7 / a8 / a6 ------------------------------------------------------------------------------------
    def find_splits(self, n_nodes, min_rows, msi):
        self.lib.call("h2ok_hist_find_splits_dev", self.ctx, _ptr(self.hist), _ptr(self.descs), n_nodes, self.cols,
                      self.max_bins, float(min_rows), float(msi), _ptr(self.splits))
        self.launches += 1

    def next_level(self, n_nodes, params, depth):
        self.lib.call("h2ok_hist_next_level_dev", self.ctx, _ptr(self.splits), _ptr(self.descs), _ptr(self.min2),
                      _ptr(self.maxin), n_nodes, self.cols, C.byref(params), depth, _ptr(self.child_descs),
                      _ptr(self.child_ids), _ptr(self.n_children))
        self.launches += 2
        return int(self.n_children.item())  # the one host sync per level: how many nodes the next level has

    def no_children(self, n_nodes):
        self.child_ids[: 2 * n_nodes].fill_(-1)

    def apply_splits(self, n_nodes, gid_base, last=False):
        self.lib.call("h2ok_hist_apply_splits_dev", self.ctx, _ptr(self.X), self.ld, self.rows, _ptr(self.splits),
                      _ptr(self.child_ids), gid_base, n_nodes, _ptr(self.nid), _ptr(self.leaf))
        self.launches += 1

    # ---- host copies (tests, tree assembly) --------------------------------------------------------------
    def splits_np(self, n):
        return self.splits[:n].cpu().numpy().view(abi.SPLIT).reshape(n).copy()

    def descs_np(self, n):
        return self.descs[:n].cpu().numpy().view(abi.COL_DESC).reshape(n, self.cols).copy()

    def stats(self):
        s = abi.Stats()
        self.lib.call("h2ok_hist_last_level_stats", self.ctx, C.byref(s))
        return s


class EngineShard(CudaShard):
    """One GPU's rows held in the library's resident tree engine (h2ok_hist_tree_*): a row-major copy of X plus row
    records kept grouped by node, so every histogram CTA works on one node at a time with its descriptors in
    registers.  Same interface and bit-identical results as CudaShard.  The column-major X only passes through a
    staging buffer of at most `stage_bytes`; it is not kept."""

    stage_bytes = 2 << 30

    def __init__(self, device=0, max_level_nodes=128, sparse=False, u8=False):
        """sparse: keep only the nonzero cells (H2OK_TREE_SPARSE) — for matrices that are mostly exact zeros.
        u8: pre-binned input, one byte per cell (H2OK_TREE_U8); the cells are integer codes 0..254, NaN / 255 = NA."""
        super().__init__(device, max_level_nodes)
        self.sparse, self.u8 = sparse, u8

    def _alloc_rows(self, rows, cols, weighted):
        self.rows, self.cols = rows, cols
        flags = (abi.TREE_WEIGHTED if weighted else 0) | (abi.TREE_SPARSE if self.sparse else 0) | (abi.TREE_U8 if self.u8 else 0)
        self.lib.call("h2ok_hist_tree_create", self.ctx, rows, cols, flags)
        n = (rows + 63) // 64 * 64
        self.ld = n
        dev = self.device
        self.X = None
        self.y = torch.empty(n, dtype=torch.float32, device=dev)
        self.w = torch.empty(n, dtype=torch.float32, device=dev) if weighted else None
        self.weighted = weighted
        self.wq = torch.empty(n, dtype=torch.int32, device=dev) if weighted else None
        self.wyq = torch.empty(n, dtype=torch.int32, device=dev)
        self.wyyq = torch.empty(n, dtype=torch.int32, device=dev)
        self._nid = torch.empty(n, dtype=torch.int32, device=dev)
        self._leaf = torch.empty(n, dtype=torch.int32, device=dev)
        self._cmn = torch.empty(cols, dtype=torch.int32, device=dev)
        self._cmx = torch.empty(cols, dtype=torch.int32, device=dev)

    def _blocks(self):
        """Column blocks whose column-major staging copy stays under stage_bytes."""
        per = max(1, min(self.cols, self.stage_bytes // (4 * self.ld)))
        return [(c0, min(per, self.cols - c0)) for c0 in range(0, self.cols, per)]

    def _load_block(self, stage, c0, nc):
        """Column-major staging block -> the engine's row-major copy; the block's column ranges are kept for a1."""
        self.lib.call("h2ok_hist_tree_load_cols_dev", self.ctx, c0, nc, _ptr(stage), self.ld)
        self.lib.call("h2ok_hist_col_ranges_dev", self.ctx, _ptr(stage), self.ld, self.rows, nc,
                      C.c_void_p(self._cmn.data_ptr() + 4 * c0), C.c_void_p(self._cmx.data_ptr() + 4 * c0))

    def set_data_synth(self, shape, seed, row0, rows, cols, na_permille=0, weight_mode=0):
        self._alloc_rows(rows, cols, weight_mode != 0)
        blocks = self._blocks()
        stage = torch.empty((blocks[0][1], self.ld), dtype=torch.float32, device=self.device)
        for c0, nc in blocks:
            self.lib.call("h2ok_hist_synth_fill_cols_dev", self.ctx, shape, seed, row0, rows, c0, nc, self.ld,
                          na_permille, _ptr(stage))
            self._load_block(stage, c0, nc)
        self.lib.call("h2ok_hist_synth_fill_dev", self.ctx, shape, seed, row0, rows, cols, self.ld, na_permille,
                      weight_mode, None, _ptr(self.y), _ptr(self.w))
        torch.cuda.current_stream(self.device).synchronize()  # the staging buffer is released here

    def set_data_synth_prebinned(self, shape, seed, row0, rows, cols, nbins=255, na_permille=0):
        """u8 shards: the generator's fp32 columns, binned ONCE against their root descriptors (global bins from the
        column ranges over all ranks, `nbins` <= 255 of them) by h2ok_hist_prebin_dev, loaded as codes.  Returns the
        root descriptors the codes were made with."""
        assert self.u8 and nbins <= 255
        self._alloc_rows(rows, cols, False)
        blocks = self._blocks()
        stage = torch.empty((blocks[0][1], self.ld), dtype=torch.float32, device=self.device)
        codes = torch.empty((blocks[0][1], self.ld), dtype=torch.uint8, device=self.device)
        params = abi.Params.make(nbins=nbins)
        descs = np.zeros(cols, abi.COL_DESC)
        for c0, nc in blocks:
            self.lib.call("h2ok_hist_synth_fill_cols_dev", self.ctx, shape, seed, row0, rows, c0, nc, self.ld, na_permille,
                          _ptr(stage))
            mn = torch.empty(nc, dtype=torch.int32, device=self.device)
            mx = torch.empty(nc, dtype=torch.int32, device=self.device)
            self.lib.call("h2ok_hist_col_ranges_dev", self.ctx, _ptr(stage), self.ld, rows, nc, _ptr(mn), _ptr(mx))
            if _world() > 1:  # every rank must bin against the same global ranges
                mn ^= _I32_SIGN
                mx ^= _I32_SIGN
                dist.all_reduce(mn, op=dist.ReduceOp.MIN)
                dist.all_reduce(mx, op=dist.ReduceOp.MAX)
                mn ^= _I32_SIGN
                mx ^= _I32_SIGN
            lo = abi.decode_f32(mn.cpu().numpy().view(np.uint32))
            hi = abi.decode_f32(mx.cpu().numpy().view(np.uint32))
            d = np.zeros(nc, abi.COL_DESC)
            self.lib.call("h2ok_hist_root_descs_from_ranges", abi.ptr(lo), abi.ptr(hi), nc, C.byref(params), None, abi.ptr(d))
            descs[c0:c0 + nc] = d
            self.lib.call("h2ok_hist_prebin_dev", self.ctx, _ptr(stage), self.ld, rows, nc, abi.ptr(d), _ptr(codes), self.ld)
            self.lib.call("h2ok_hist_tree_load_bins_dev", self.ctx, c0, nc, _ptr(codes), self.ld)
        self.lib.call("h2ok_hist_synth_fill_dev", self.ctx, shape, seed, row0, rows, cols, self.ld, 0, 0, None, _ptr(self.y), None)
        torch.cuda.current_stream(self.device).synchronize()
        return descs

    def set_data_tensors(self, X, y, w=None):
        cols, rows = X.shape
        self._alloc_rows(rows, cols, w is not None)
        blocks = self._blocks()
        stage = torch.zeros((blocks[0][1], self.ld), dtype=torch.float32, device=self.device)
        for c0, nc in blocks:
            stage[:nc, :rows].copy_(X[c0:c0 + nc], non_blocking=False)
            self._load_block(stage, c0, nc)
        self.set_targets(y, w)
        torch.cuda.current_stream(self.device).synchronize()

    def col_ranges(self):
        return self._cmn.clone(), self._cmx.clone()

    @property
    def nid(self):
        self.lib.call("h2ok_hist_tree_nid_dev", self.ctx, _ptr(self._nid))
        return self._nid

    @property
    def leaf(self):
        self.lib.call("h2ok_hist_tree_leaf_dev", self.ctx, _ptr(self._leaf))
        return self._leaf

    def quantize(self, quant):
        super().quantize(quant)
        self.lib.call("h2ok_hist_tree_begin_dev", self.ctx, _ptr(self.wq), _ptr(self.wyq), _ptr(self.wyyq), C.byref(quant))
        self.launches += 1

    def reset_nodes(self):
        pass  # tree_begin_dev put every row back at the root

    def build_level(self, n_nodes, level_bins=0):
        if level_bins:
            self.lib.call("h2ok_hist_tree_level_bins", self.ctx, int(level_bins))
        self.lib.call("h2ok_hist_tree_build_level_dev", self.ctx, n_nodes, _ptr(self.descs), self.max_bins, 1,
                      _ptr(self.acc), _ptr(self.mn), _ptr(self.mx))
        self.launches += 2  # init_enc + histogram pass
        return self.acc[:n_nodes], self.mn[:n_nodes], self.mx[:n_nodes]

    def apply_splits(self, n_nodes, gid_base, last=False):
        self.lib.call("h2ok_hist_tree_apply_splits_dev", self.ctx, _ptr(self.splits), _ptr(self.child_ids), gid_base,
                      n_nodes, 1 if last else 0)
        self.launches += 2 if last else 3  # route, scan (+ scatter unless the tree ends here)


class TreeLoop:
    """The level loop that runs INSIDE libh2ok_hist.so (h2ok_hist_loop_*): one C call per tree level, the cross-rank
    reduce done by the library over NCCL.  Python only hands over pointers; torch.distributed is used for one thing,
    carrying the 128-byte rendezvous id from rank 0 to the other ranks when the process group exists.

    shard: an EngineShard whose data is loaded.  Same results as LevelDriver, bit for bit."""

    def __init__(self, shard, params, is_int=None, reduce_mode=abi.REDUCE_SCATTER, checksum=False, time_phases=False,
                 forest=1):
        """forest = K > 1: K trees on the same rows built in lockstep (the K class trees of a multinomial iteration) —
        sparse shards only; start them with begin_forest(ys)."""
        self.shard, self.params, self.lib = shard, params, shard.lib
        self.forest = int(forest)
        if is_int is None and getattr(shard, "u8", False):
            is_int = np.ones(shard.cols, np.int32)  # the cells of a u8 shard are integer codes
        self.is_int = None if is_int is None else np.ascontiguousarray(is_int, np.int32)
        self.max_bins = max(params.nbins, params.nbins_top_level)
        if self.forest > 1:
            self.lib.call("h2ok_hist_loop_create_forest", shard.ctx, C.byref(params), abi.ptr(self.is_int), self.forest)
        else:
            self.lib.call("h2ok_hist_loop_create", shard.ctx, C.byref(params), abi.ptr(self.is_int))
        self.lib.call("h2ok_hist_loop_set_reduce", shard.ctx, reduce_mode)
        self.time_phases = bool(time_phases)
        self.set_options(checksum=checksum)
        self.n_nodes = 0
        self.depth = 0

    def set_options(self, checksum=None, time_phases=None):
        """Switch the result checksum (three hash kernels per level) / the per-phase event timing on or off; takes
        effect with the next tree."""
        if checksum is not None:
            self.checksum_on = bool(checksum)
        if time_phases is not None:
            self.time_phases = bool(time_phases)
        self.lib.call("h2ok_hist_loop_set_options", self.shard.ctx, 1 if self.checksum_on else 0, 1 if self.time_phases else 0)

    @staticmethod
    def join_ranks(shard):
        """Every rank's context joins one NCCL communicator (call before the data is loaded / TreeLoop is made)."""
        world = _world()
        if world == 1:
            return
        ident = torch.zeros(abi.COMM_ID_BYTES, dtype=torch.uint8)
        if dist.get_rank() == 0:
            shard.lib.call("h2ok_hist_comm_unique_id", C.c_void_p(ident.data_ptr()))
        if dist.get_backend() == "nccl":
            dev = ident.to(shard.device)
            dist.broadcast(dev, 0)
            ident = dev.cpu()
        else:
            dist.broadcast(ident, 0)
        shard.lib.call("h2ok_hist_comm_init", shard.ctx, C.c_void_p(ident.data_ptr()), dist.get_rank(), world)

    def root_descs(self):
        out = np.zeros(self.shard.cols, abi.COL_DESC)
        self.lib.call("h2ok_hist_loop_root_descs", self.shard.ctx, abi.ptr(out))
        return out

    def begin_tree(self, y=None, w=None):
        """y / w: device tensors in original row order (default: the shard's own targets)."""
        y = self.shard.y if y is None else y
        w = (self.shard.w if self.shard.weighted else None) if y is self.shard.y and w is None else w
        self.lib.call("h2ok_hist_loop_begin_tree_dev", self.shard.ctx, _ptr(y), _ptr(w))
        self.shard.launches += 5  # absmax, quantise, root mean, begin, + the library's own memsets do not count
        self.n_nodes, self.depth = 1, 0

    def begin_forest(self, ys):
        """ys: device tensor [forest, ld] — the trees' target vectors in original row order."""
        assert ys.shape[0] == self.forest and ys.stride(1) == 1
        self.lib.call("h2ok_hist_loop_begin_forest_dev", self.shard.ctx, _ptr(ys), ys.stride(0))
        self.shard.launches += 5 * self.forest + 1
        self.n_nodes, self.depth = self.forest, 0

    def forest_quants(self):
        """The fixed-point scales of the current forest's trees (list of abi.Quant)."""
        out = (abi.Quant * self.forest)()
        self.lib.call("h2ok_hist_loop_forest_quants", self.shard.ctx, out)
        return list(out)

    def begin_tree_host(self, y_host, w_host=None):
        """y_host / w_host: host tensors or numpy arrays (fp32); the library copies them in."""
        p = lambda a: None if a is None else C.c_void_p(a.data_ptr() if hasattr(a, "data_ptr") else a.ctypes.data)
        self.lib.call("h2ok_hist_loop_begin_tree_host", self.shard.ctx, p(y_host), p(w_host))
        self.shard.launches += 5
        self.n_nodes, self.depth = 1, 0

    def level(self, out=None):
        done, nch = C.c_int32(), C.c_int32()
        self.lib.call("h2ok_hist_loop_level", self.shard.ctx, None if out is None else C.byref(out), C.byref(done),
                      C.byref(nch))
        last = self.depth + 1 >= self.params.max_depth
        # init_enc, histogram pass, split points, bestCol, [children: descs + compact], route, scan, [scatter]
        self.shard.launches += (6 if last else 9) + 2 * (self.forest - 1)  # a forest: totals + routing per class tree
        self.depth += 1
        self.n_nodes = nch.value
        return done.value, nch.value

    def n_nodes_left(self):
        return self.n_nodes if self.depth < self.params.max_depth else 0

    def train_tree(self, y=None, w=None, after_level=None):
        self.begin_tree(y, w)
        levels = 0
        while self.n_nodes_left() > 0:
            self.level()
            levels += 1
            if after_level:
                after_level(self)
        return levels

    def view(self):
        v = abi.LoopView()
        self.lib.call("h2ok_hist_loop_view", self.shard.ctx, C.byref(v))
        return v

    def phase_stats(self, reset=False):
        """Per tree depth: (levels timed, nodes, [hist kernel, clear + hist, reduce, search + children, routing] ms summed)."""
        d = int(self.params.max_depth)
        cnt, nodes, ms = np.zeros(d, np.int32), np.zeros(d, np.int32), np.zeros((d, 5), np.float32)
        self.lib.call("h2ok_hist_loop_phase_stats", self.shard.ctx, 1 if reset else 0, d, abi.ptr(cnt), abi.ptr(nodes), abi.ptr(ms))
        return cnt, nodes, ms

    def reset_phase_stats(self):
        self.lib.call("h2ok_hist_loop_phase_stats", self.shard.ctx, 1, 0, None, None, None)

    def checksum(self, row0=0):
        out = (C.c_uint64 * 3)()
        self.lib.call("h2ok_hist_loop_checksum", self.shard.ctx, int(row0), out)
        return [int(x) for x in out]

    # ---- host copies of the level that ran last (tests) ------------------------------------------------------
    def _dev_array(self, ptr, nbytes):
        out = np.empty(nbytes, np.uint8)
        torch.cuda.current_stream(self.shard.device).synchronize()
        self.lib.call("h2ok_hist_copy_to_host", self.shard.ctx, C.c_void_p(ptr), abi.ptr(out), nbytes)
        return out

    def done_level(self):
        """dict of numpy copies: descs, acc, min2, maxin (decoded), splits, child_ids, n, nch, pair0, pair1."""
        v = self.view()
        n, cols, nb1 = v.done_nodes, v.cols, v.max_bins + 1
        out = dict(n=n, nch=v.done_children, pair0=v.done_pair0, pair1=v.done_pair1, depth=v.done_depth,
                   gid_base=v.done_gid_base, quant=v.quant)
        out["descs"] = self._dev_array(v.d_done_descs, n * cols * abi.COL_DESC.itemsize).view(abi.COL_DESC).reshape(n, cols)
        out["acc"] = self._dev_array(v.d_done_acc, n * cols * nb1 * 3 * 8).view(np.int64).reshape(n, cols, nb1, 3)
        mn = self._dev_array(v.d_done_min2_enc, n * cols * 4).view(np.uint32).reshape(n, cols)
        mx = self._dev_array(v.d_done_maxin_enc, n * cols * 4).view(np.uint32).reshape(n, cols)
        out["min2"] = abi.decode_f32(mn) + np.float32(0)
        out["maxin"] = abi.decode_f32(mx) + np.float32(0)
        out["splits"] = self._dev_array(v.d_done_splits, n * abi.SPLIT.itemsize).view(abi.SPLIT).reshape(n)
        out["child_ids"] = self._dev_array(v.d_done_child_ids, 2 * n * 4).view(np.int32)
        if v.done_children:
            out["child_descs"] = self._dev_array(v.d_descs, v.done_children * cols * abi.COL_DESC.itemsize).view(
                abi.COL_DESC).reshape(v.done_children, cols)
        return out


def _world():
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def reduce_level(acc, mn, mx, wait=True):
    """a5 — MRTask.reduce / DHistogram.add across ranks: integer sum of the fixed-point histogram, integer min / max
    of the encoded observed ranges.  The encodings are unsigned; flipping the top bit makes signed compares agree,
    and min(a) = ~max(~a) folds the two range reductions into ONE max all-reduce.  With wait=False the range
    reduction is left in flight (only next_level needs it, the split search does not) and the returned function
    completes it."""
    if _world() == 1:
        return lambda: None
    dist.all_reduce(acc, op=dist.ReduceOp.SUM)
    n = mn.numel()
    buf = torch.empty(2 * n, dtype=mn.dtype, device=mn.device)
    # ~(x ^ SIGN) == x ^ 0x7fffffff: one xor per half on the way in, one on the way out
    torch.bitwise_xor(mn.reshape(-1), _I32_MAX, out=buf[:n])
    torch.bitwise_xor(mx.reshape(-1), _I32_SIGN, out=buf[n:])
    work = dist.all_reduce(buf, op=dist.ReduceOp.MAX, async_op=True)

    def finish():
        work.wait()
        torch.bitwise_xor(buf[:n], _I32_MAX, out=mn.view(-1))
        torch.bitwise_xor(buf[n:], _I32_SIGN, out=mx.view(-1))

    if wait:
        finish()
        return lambda: None
    return finish


class LevelDriver:
    """SharedTree.buildLayer loop over one shard per rank.  `shard` is a CudaShard (tests substitute a CPU double
    with the same methods to exercise the reduce logic under gloo)."""

    def __init__(self, shard, params, is_int=None, qbits=25):
        self.shard = shard
        self.params = params
        self.is_int = None if is_int is None else np.ascontiguousarray(is_int, np.int32)
        self.qbits = qbits
        self.lib = shard.lib
        self.max_bins = max(params.nbins, params.nbins_top_level)
        self.root = None
        self.n_nodes = 0
        self.depth = 0
        self.gid_base = 0

    # -- once per data set: root descriptors from the global column ranges
    def prepare_data(self):
        mn, mx = self.shard.col_ranges()
        if _world() > 1:
            mn ^= _I32_SIGN
            mx ^= _I32_SIGN
            dist.all_reduce(mn, op=dist.ReduceOp.MIN)
            dist.all_reduce(mx, op=dist.ReduceOp.MAX)
            mn ^= _I32_SIGN
            mx ^= _I32_SIGN
        lo = abi.decode_f32(mn.cpu().numpy().view(np.uint32))
        hi = abi.decode_f32(mx.cpu().numpy().view(np.uint32))
        out = np.zeros(self.shard.cols, abi.COL_DESC)
        self.lib.call("h2ok_hist_root_descs_from_ranges", abi.ptr(lo), abi.ptr(hi), self.shard.cols,
                      C.byref(self.params), abi.ptr(self.is_int), abi.ptr(out))
        self.root = out
        max_nodes = min(2 ** max(self.params.max_depth - 1, 0), 1023)  # H2OK_MAX_LEVEL_NODES
        self.shard.ensure_level_buffers(max_nodes, self.max_bins)
        return out

    # -- once per tree: common fixed-point scales, quantised targets, all rows at the root
    def prepare_tree(self):
        m = self.shard.target_absmax()
        if _world() > 1:
            dist.all_reduce(m, op=dist.ReduceOp.MAX)
        q = abi.Quant()
        mh = m.cpu().numpy()
        self.lib.call("h2ok_hist_choose_quant", abi.ptr(mh), 1 if self.shard.weighted else 0, self.qbits, C.byref(q))
        self.shard.quantize(q)
        self.shard.reset_nodes()
        self.n_nodes = self.shard.upload_descs(self.root.reshape(1, -1))
        self.gid_base = 0
        self.depth = 0
        return q

    # -- one level = one "step" of the benchmark
    def level(self, last=None, inspect=None):
        """Build, reduce, search, derive children, route.  `inspect(driver, n_nodes, n_children)` runs after the
        children are known and before the row map and descriptors are overwritten (tests read the level there)."""
        sh, p = self.shard, self.params
        n = self.n_nodes
        # H2O halves nbins_top_level per level (Split.nextLevelHistos: max(parent bins >> 1, nbins))
        acc, mn, mx = sh.build_level(n, level_bins=max(p.nbins, p.nbins_top_level >> self.depth))
        finish_ranges = reduce_level(acc, mn, mx, wait=False)
        sh.finalize_level(n, ranges=False)  # the split search starts while the range reduction is still in flight
        sh.find_splits(n, p.min_rows, p.min_split_improvement)
        finish_ranges()
        sh.finalize_level(n, hist=False)
        last = (self.depth + 1 >= p.max_depth) if last is None else last
        if last:
            sh.no_children(n)
            nch = 0
        else:
            nch = sh.next_level(n, p, self.depth)
        if inspect:
            inspect(self, n, nch)
        sh.apply_splits(n, self.gid_base, last=(nch == 0))
        self.gid_base += n
        self.depth += 1
        done_nodes = n
        if nch > 0:
            sh.adopt_children(nch)
        self.n_nodes = nch
        return done_nodes, nch

    def n_nodes_left(self):
        """Active nodes waiting for the next level; 0 when the current tree is finished (or none was started)."""
        return self.n_nodes if self.depth < self.params.max_depth else 0

    def train_tree(self, inspect=None):
        self.prepare_tree()
        levels = 0
        while self.n_nodes > 0 and self.depth < self.params.max_depth:
            self.level(inspect=inspect)
            levels += 1
        return levels
