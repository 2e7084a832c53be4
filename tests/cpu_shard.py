"""CPU test double for h2ok_hist.driver.CudaShard, backed by the oracle.  TEST INFRASTRUCTURE ONLY: it lets the
host-side LevelDriver logic (sharding, the cross-rank reduce of the integer histograms and encoded ranges, level
bookkeeping) run under the gloo backend on a machine without GPUs.  The product never imports this file."""
import ctypes as C

import numpy as np
import torch

from h2ok_hist import abi, capi


class CpuShard:
    def __init__(self, oracle, X, rows, y, w=None):
        self.oracle = oracle
        self.lib = capi.load()  # only its pure-host helpers are used (root descriptors, scale choice)
        self.device = torch.device("cpu")
        self.X, self.rows, self.cols = X, rows, X.shape[0]
        self.y, self.w = y, w
        self.weighted = w is not None
        self.launches = 0

    def col_ranges(self):
        mn, mx = self.oracle.col_ranges(self.X, self.rows)
        enc = lambda a, empty: torch.from_numpy(abi.encode_f32(a).view(np.int32).copy())
        if self.rows == 0:
            mn[:] = np.inf
            mx[:] = -np.inf
        return enc(mn, None), enc(mx, None)

    def target_absmax(self):
        m = np.zeros(3)
        self.oracle.lib.h2ok_oracle_target_absmax(abi.ptr(self.y), abi.ptr(self.w), self.rows, abi.ptr(m))
        return torch.from_numpy(m)

    def quantize(self, q):
        self.quant = q
        self.wq, self.wyq, self.wyyq = self.oracle.quantize_rows(self.y, self.w, q)

    def reset_nodes(self):
        self.nid = np.zeros(max(self.rows, 1), np.int32)
        self.leaf = np.full(max(self.rows, 1), -1, np.int32)

    def ensure_level_buffers(self, max_nodes, max_bins):
        self.max_bins = max_bins

    def upload_descs(self, d):
        self.descs = np.ascontiguousarray(d, abi.COL_DESC).reshape(-1, self.cols).copy()
        return self.descs.shape[0]

    def adopt_children(self, n):
        self.descs = self.child[:n].copy()

    def build_level(self, n, level_bins=0):
        acc, min2, maxin = self.oracle.build_level_q(self.X, self.rows, self.nid, self.wq, self.wyq, self.wyyq,
                                                     self.descs[:n], self.max_bins)
        self.acc = torch.from_numpy(acc)
        self.mn = torch.from_numpy(abi.encode_f32(min2).view(np.int32).copy())
        self.mx = torch.from_numpy(abi.encode_f32(maxin).view(np.int32).copy())
        return self.acc, self.mn, self.mx

    def finalize_level(self, n, hist=True, ranges=True):
        if hist:
            self.hist = self.oracle.finalize(self.acc.numpy(), self.quant, self.max_bins)
        if ranges:
            self.min2 = abi.decode_f32(self.mn.numpy().view(np.uint32)) + np.float32(0)
            self.maxin = abi.decode_f32(self.mx.numpy().view(np.uint32)) + np.float32(0)

    def find_splits(self, n, min_rows, msi):
        self.splits = self.oracle.find_splits(self.hist, self.descs[:n], min_rows, msi)

    def next_level(self, n, params, depth):
        self.child, self.child_ids, nch = self.oracle.next_level(self.splits, self.descs[:n], self.min2, self.maxin, params, depth)
        return nch

    def no_children(self, n):
        self.child_ids = np.full(2 * n, -1, np.int32)

    def apply_splits(self, n, gid_base, last=False):
        self.oracle.apply_splits(self.X, self.rows, self.splits, self.child_ids, gid_base, self.nid, self.leaf)
