"""a9 — the tree-level caller of the histogram path: GBM.buildNextKTrees for one-class
trees (gaussian, bernoulli) as H2O-3 runs it [hex/tree/gbm/GBM.java, recalled]: residual pass → one tree through
LevelDriver → GammaPass → AddTreeContributions.  All arithmetic is in libh2ok_hist.so; this file only sequences it and
does the one cross-rank reduce of the leaf sums (integers, so every rank derives the same leaf values).
"""
import ctypes as C
import math

import numpy as np
import torch
import torch.distributed as dist_

from . import abi
from .driver import _ptr, _world


def leaf_gamma(num, den, scale_num, scale_den):
    """Leaf values from the integer GammaPass sums: (num / scale_num) / (den / scale_den), 0 for empty leaves."""
    num = num.astype(np.float64) / scale_num
    den = den.astype(np.float64) / scale_den
    return np.where(den != 0, num / np.where(den != 0, den, 1.0), 0.0)


class GBMTrainer:
    def __init__(self, driver, label, dist=abi.DIST_GAUSSIAN, learn_rate=0.1):
        """driver: a LevelDriver whose shard already holds X (prepare_data() done); label: device tensor [rows]."""
        self.drv, self.shard = driver, driver.shard
        self.lib = driver.lib
        self.dist, self.lr = dist, float(learn_rate)
        self.rows = self.shard.rows
        self.label = label
        self.f = torch.zeros(self.rows, dtype=torch.float32, device=self.shard.device)
        self.denq = torch.zeros(self.rows, dtype=torch.int32, device=self.shard.device) if dist == abi.DIST_BERNOULLI else None
        self.trees = []

    def init_f(self):
        """H2O's _initial_f: the (weighted) mean response, on the link scale."""
        w = self.shard.w[: self.rows].double() if self.shard.weighted else None
        y = self.label.double()
        s = torch.stack([(y * w).sum() if w is not None else y.sum(), w.sum() if w is not None else torch.tensor(float(self.rows), device=y.device, dtype=torch.float64)])
        if _world() > 1:
            dist_.all_reduce(s)
        mean = float(s[0] / s[1])
        f0 = mean if self.dist == abi.DIST_GAUSSIAN else math.log(mean / (1.0 - mean))
        self.f.fill_(np.float32(f0))
        return float(np.float32(f0))

    def step(self):
        """One boosting iteration; returns the tree's leaf values indexed by leaf code."""
        sh, rows = self.shard, self.rows
        den_scale = 2.0**28 if self.dist == abi.DIST_BERNOULLI else 1.0
        self.lib.call("h2ok_hist_gbm_residual_dev", sh.ctx, self.dist, _ptr(self.label), _ptr(self.f), _ptr(sh.w), rows,
                      den_scale, _ptr(sh.y), _ptr(self.denq))
        levels = []

        def keep(d, n, nch):
            levels.append((d.shard.splits_np(n), d.shard.child_ids[: 2 * n].cpu().numpy().copy()))

        self.drv.train_tree(inspect=keep)
        n_codes = 3 * self.drv.gid_base
        q = sh.quant
        if self.dist == abi.DIST_GAUSSIAN:
            denq, den_scale = (sh.wq, q.scale_w) if sh.weighted else (None, 1.0)
        else:
            denq = self.denq
        num = torch.zeros(n_codes, dtype=torch.int64, device=sh.device)
        den = torch.zeros(n_codes, dtype=torch.int64, device=sh.device)
        leaf = sh.leaf
        self.lib.call("h2ok_hist_gbm_gamma_pass_dev", sh.ctx, _ptr(leaf), _ptr(sh.wyq), _ptr(denq), rows, n_codes,
                      _ptr(num), _ptr(den))
        if _world() > 1:
            dist_.all_reduce(num)
            dist_.all_reduce(den)
        gamma = leaf_gamma(num.cpu().numpy(), den.cpu().numpy(), q.scale_wy, den_scale)
        dg = torch.from_numpy(gamma).to(sh.device)
        self.lib.call("h2ok_hist_gbm_add_gamma_dev", sh.ctx, _ptr(leaf), rows, _ptr(dg), n_codes, self.lr, _ptr(self.f))
        sh.launches += 3
        self.trees.append(dict(levels=levels, gamma=gamma))
        return gamma

    def train(self, n_trees):
        for _ in range(n_trees):
            self.step()
        return self.f

    def loss(self):
        """Training loss (MSE or logloss) of the current f — reporting only."""
        y, f = self.label.double(), self.f.double()
        if self.dist == abi.DIST_GAUSSIAN:
            s = torch.stack([((y - f) ** 2).sum(), torch.tensor(float(self.rows), device=y.device, dtype=torch.float64)])
        else:
            p = torch.sigmoid(f).clamp(1e-15, 1 - 1e-15)
            s = torch.stack([-(y * p.log() + (1 - y) * (1 - p).log()).sum(), torch.tensor(float(self.rows), device=y.device, dtype=torch.float64)])
        if _world() > 1:
            dist_.all_reduce(s)
        return float(s[0] / s[1])
