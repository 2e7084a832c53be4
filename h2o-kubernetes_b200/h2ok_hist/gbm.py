"""a9 — the tree-level caller of the histogram path: GBM.buildNextKTrees as H2O-3 runs it
[hex/tree/gbm/GBM.java, recalled] for one-class trees (gaussian, bernoulli) and for K-class multinomial (K trees per
iteration on softmax residuals, the workload of BASELINE config 4): residual pass → tree(s) through LevelDriver →
GammaPass → AddTreeContributions.  All arithmetic is in libh2ok_hist.so; this file only sequences it and
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
    def __init__(self, driver, label, dist=abi.DIST_GAUSSIAN, learn_rate=0.1, n_classes=None, keep_levels=True):
        """driver: a LevelDriver (prepare_data() done) or a TreeLoop (the level loop inside the library) whose shard
        already holds X; label: device tensor [rows] (multinomial: class ids 0..n_classes-1 as floats).
        keep_levels: copy every level's splits and child ids to the host (tests); off for timing runs."""
        self.drv, self.shard = driver, driver.shard
        self.lib = driver.lib
        self.in_library = hasattr(driver, "done_level")
        self.keep_levels = keep_levels
        self.dist, self.lr = dist, float(learn_rate)
        self.rows = self.shard.rows
        self.label = label
        dev = self.shard.device
        self.trees = []
        if dist == abi.DIST_MULTINOMIAL:
            if n_classes is None or n_classes < 2:
                raise ValueError("multinomial needs n_classes >= 2")
            self.K = int(n_classes)
            self.ld = (self.rows + 63) // 64 * 64
            self.f = torch.zeros((self.K, self.ld), dtype=torch.float32, device=dev)
            self.res = torch.zeros((self.K, self.ld), dtype=torch.float32, device=dev)
            self.denq = torch.zeros((self.K, self.ld), dtype=torch.int32, device=dev)
            return
        self.K = 1
        self.f = torch.zeros(self.rows, dtype=torch.float32, device=dev)
        self.denq = torch.zeros(self.rows, dtype=torch.int32, device=dev) if dist == abi.DIST_BERNOULLI else None

    def init_f(self):
        """H2O's _initial_f: the (weighted) mean response, on the link scale; multinomial starts from zero scores."""
        if self.dist == abi.DIST_MULTINOMIAL:
            self.f.zero_()
            return 0.0
        w = self.shard.w[: self.rows].double() if self.shard.weighted else None
        y = self.label.double()
        s = torch.stack([(y * w).sum() if w is not None else y.sum(), w.sum() if w is not None else torch.tensor(float(self.rows), device=y.device, dtype=torch.float64)])
        if _world() > 1:
            dist_.all_reduce(s)
        mean = float(s[0] / s[1])
        f0 = mean if self.dist == abi.DIST_GAUSSIAN else math.log(mean / (1.0 - mean))
        self.f.fill_(np.float32(f0))
        return float(np.float32(f0))

    def _tree_and_gamma(self, denq, den_scale, f_out, factor=1.0):
        """One tree on the targets in shard.y, its GammaPass and AddTreeContributions into f_out."""
        sh, rows = self.shard, self.rows
        levels = []
        if self.in_library:  # one C call per level; the quantised targets live inside the context
            def after(loop):
                if self.keep_levels:
                    lv = loop.done_level()
                    levels.append((lv["splits"], lv["child_ids"]))

            self.drv.train_tree(after_level=after)
            v = self.drv.view()
            n_codes, q = 3 * v.gid_base, v.quant
            p_wyq, p_wq = C.c_void_p(v.d_wyq), C.c_void_p(v.d_wq) if v.d_wq else None
        else:
            def keep(d, n, nch):
                if self.keep_levels:
                    levels.append((d.shard.splits_np(n), d.shard.child_ids[: 2 * n].cpu().numpy().copy()))

            self.drv.train_tree(inspect=keep)
            n_codes, q = 3 * self.drv.gid_base, sh.quant
            p_wyq, p_wq = _ptr(sh.wyq), _ptr(sh.wq)
        if den_scale is None:  # gaussian
            p_den, den_scale = (p_wq, q.scale_w) if sh.weighted else (None, 1.0)
        else:
            p_den = _ptr(denq)
        num = torch.zeros(n_codes, dtype=torch.int64, device=sh.device)
        den = torch.zeros(n_codes, dtype=torch.int64, device=sh.device)
        leaf = sh.leaf
        self.lib.call("h2ok_hist_gbm_gamma_pass_dev", sh.ctx, _ptr(leaf), p_wyq, p_den, rows, n_codes, _ptr(num), _ptr(den))
        if _world() > 1:
            dist_.all_reduce(num)
            dist_.all_reduce(den)
        gamma = factor * leaf_gamma(num.cpu().numpy(), den.cpu().numpy(), q.scale_wy, den_scale)
        dg = torch.from_numpy(gamma).to(sh.device)
        self.lib.call("h2ok_hist_gbm_add_gamma_dev", sh.ctx, _ptr(leaf), rows, _ptr(dg), n_codes, self.lr, _ptr(f_out))
        sh.launches += 2
        return dict(levels=levels, gamma=gamma)

    def _step_multinomial_forest(self):
        """The K class trees of the iteration as ONE forest (h2ok_hist_loop_create_forest): every level streams the
        sparse column lists once for all K trees.  Same trees, leaf values and scores as _step_multinomial."""
        sh, rows, K = self.shard, self.rows, self.K
        den_scale = 2.0**28
        self.lib.call("h2ok_hist_gbm_residual_multi_dev", sh.ctx, K, _ptr(self.label), _ptr(self.f), self.ld, _ptr(sh.w),
                      rows, den_scale, _ptr(self.res), _ptr(self.denq))
        sh.launches += 1
        self.drv.begin_forest(self.res)
        while self.drv.n_nodes_left() > 0:
            self.drv.level()
        v = self.drv.view()
        quants = self.drv.forest_quants()
        n_codes = 3 * v.gid_base
        pad = (rows + 3) // 4 * 4
        out = []
        nums = torch.zeros((K, n_codes), dtype=torch.int64, device=sh.device)
        dens = torch.zeros((K, n_codes), dtype=torch.int64, device=sh.device)
        for k in range(K):
            self.lib.call("h2ok_hist_gbm_gamma_pass_dev", sh.ctx, C.c_void_p(v.d_leaf + 4 * k * rows),
                          C.c_void_p(v.d_wyq + 4 * k * pad), _ptr(self.denq[k]), rows, n_codes, _ptr(nums[k]), _ptr(dens[k]))
        if _world() > 1:
            dist_.all_reduce(nums)
            dist_.all_reduce(dens)
        nums_h, dens_h = nums.cpu().numpy(), dens.cpu().numpy()
        for k in range(K):
            gamma = (K - 1.0) / K * leaf_gamma(nums_h[k], dens_h[k], quants[k].scale_wy, den_scale)
            dg = torch.from_numpy(gamma).to(sh.device)
            self.lib.call("h2ok_hist_gbm_add_gamma_dev", sh.ctx, C.c_void_p(v.d_leaf + 4 * k * rows), rows, _ptr(dg), n_codes,
                          self.lr, _ptr(self.f[k]))
            sh.launches += 2
            self.trees.append(dict(levels=[], gamma=gamma, cls=k))
            out.append(gamma)
        return out

    def _step_multinomial(self):
        """K trees on the softmax residuals of the current scores (all residuals are taken before the first tree)."""
        if getattr(self.drv, "forest", 1) == self.K and self.K > 1:
            return self._step_multinomial_forest()
        sh, rows, K = self.shard, self.rows, self.K
        den_scale = 2.0**28
        self.lib.call("h2ok_hist_gbm_residual_multi_dev", sh.ctx, K, _ptr(self.label), _ptr(self.f), self.ld, _ptr(sh.w),
                      rows, den_scale, _ptr(self.res), _ptr(self.denq))
        sh.launches += 1
        out = []
        for k in range(K):
            sh.y[:rows].copy_(self.res[k, :rows])
            t = self._tree_and_gamma(self.denq[k], den_scale, self.f[k], factor=(K - 1.0) / K)
            t["cls"] = k
            self.trees.append(t)
            out.append(t["gamma"])
        return out

    def step(self):
        """One boosting iteration; returns the tree's leaf values indexed by leaf code (multinomial: one per class)."""
        if self.dist == abi.DIST_MULTINOMIAL:
            return self._step_multinomial()
        sh, rows = self.shard, self.rows
        den_scale = 2.0**28 if self.dist == abi.DIST_BERNOULLI else 1.0
        self.lib.call("h2ok_hist_gbm_residual_dev", sh.ctx, self.dist, _ptr(self.label), _ptr(self.f), _ptr(sh.w), rows,
                      den_scale, _ptr(sh.y), _ptr(self.denq))
        sh.launches += 1
        if self.dist == abi.DIST_GAUSSIAN:
            # the denominator is the weight (quantised with the histogram's own scale, known after the tree's
            # prepare_tree) or the row count: resolved lazily below
            t = self._tree_and_gamma(None, None, self.f)
        else:
            t = self._tree_and_gamma(self.denq, den_scale, self.f)
        self.trees.append(t)
        return t["gamma"]

    def train(self, n_trees):
        for _ in range(n_trees):
            self.step()
        return self.f

    def loss(self):
        """Training loss (MSE or logloss) of the current f — reporting only."""
        y = self.label.double()
        f = self.f.double() if self.dist != abi.DIST_MULTINOMIAL else None
        if self.dist == abi.DIST_MULTINOMIAL:
            logp = torch.log_softmax(self.f[:, : self.rows].double(), dim=0)
            picked = logp.gather(0, self.label.long()[None, :])[0]
            s = torch.stack([-picked.sum(), torch.tensor(float(self.rows), device=y.device, dtype=torch.float64)])
        elif self.dist == abi.DIST_GAUSSIAN:
            s = torch.stack([((y - f) ** 2).sum(), torch.tensor(float(self.rows), device=y.device, dtype=torch.float64)])
        else:
            p = torch.sigmoid(f).clamp(1e-15, 1 - 1e-15)
            s = torch.stack([-(y * p.log() + (1 - y) * (1 - p).log()).sum(), torch.tensor(float(self.rows), device=y.device, dtype=torch.float64)])
        if _world() > 1:
            dist_.all_reduce(s)
        return float(s[0] / s[1])
