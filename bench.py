#!/usr/bin/env python
"""bench.py — GBM histogram rows*cols/s per tree level (BASELINE.json metric) on N B200s.

A "step" is one tree level of the hot path on the 100M x 64, 256-bin, depth-8 workload (BASELINE config 3, the
configuration the metric is quoted on): score_decide of the previous level's splits → histogram build (the
dominant kernel) → cross-rank reduce (N>1) → split search → children descriptors.  Steps cycle through levels
0..7 of real trees, so node counts and the row→node map are the ones the algorithm produces.  Rows are sharded
across ranks (strong scaling: total rows fixed); launch under torchrun for N>1.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for the definition of every field.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "h2o-kubernetes_b200"), os.path.join(ROOT, "tests"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "gbm_histogram_rows_x_cols_per_s_per_tree_level"
UNIT = "cells/s"
SEED = 20261106
WORKLOADS = {
    "c3": dict(name="GBM histogram build 100M rows x 64 cols, 256 bins, depth 8 (BASELINE config 3)", rows=100_000_000,
               cols=64, bins=256, depth=8, shape=3),
    "c2": dict(name="GBM histogram build 1M rows x 28 cols (HIGGS-shaped), 256 bins, depth 6 (BASELINE config 2)",
               rows=1_000_000, cols=28, bins=256, depth=6, shape=2),
    # config 4: the K = 10 class trees of one multinomial iteration, one after the other on the sparse storage (a step is
    # one class-tree level; ten steps are one "fused level" of SURVEY.md §8(d), reported as config.fused_level_ms)
    "c4": dict(name="GBM histogram build 10M rows x 2000 sparse-ish cols, 64 bins, depth 6, multinomial 10-class targets "
                    "(BASELINE config 4), class trees one after the other", rows=10_000_000, cols=2000, bins=64, depth=6,
               shape=4, k=10, sparse=True),
    # the same ten trees built in lockstep as a forest: a step is one level of ALL ten trees, value counts the ten
    # class-tree levels in it.  Measured slower than "c4" (DESIGN.md §3.4), kept as the comparison.
    "c4-forest": dict(name="GBM histogram build 10M rows x 2000 sparse-ish cols, 64 bins, depth 6, multinomial 10-class "
                           "targets (BASELINE config 4), the ten class trees as one forest", rows=10_000_000, cols=2000,
                      bins=64, depth=6, shape=4, k=10, sparse=True, forest=True),
    # config 3's matrix binned ONCE against its root descriptors (255 global bins) and kept as one byte per cell
    "c3-u8": dict(name="GBM histogram build 100M rows x 64 cols pre-binned to u8 codes (255 global bins + NA), depth 8 "
                       "(BASELINE config 3's matrix, input format of SURVEY.md §8(f) row 4)", rows=100_000_000, cols=64, bins=255,
                  depth=8, shape=3, u8=True),
    "c4-dense": dict(name="GBM histogram build 10M rows x 2000 sparse-ish cols, 64 bins, depth 6, 10-class targets, DENSE "
                          "fp32 storage (BASELINE config 4 as round 1 ran it)", rows=10_000_000, cols=2000, bins=64, depth=6,
                     shape=4, k=10),
}
CPU_SAMPLE_ROWS = 2_000_000
CPU_SAMPLE_CELLS = 128_000_000  # and at most this many cells per level (2M rows x 64 columns; 64K rows x 2000)


def algo_bytes_level(rows, cols, n_nodes, bins, weighted=False, sparse_nnz=None, u8=False, trees=1):
    """every fp32 cell once, nid + y (+ w) once per row, the fp64 histogram written once.
    Sparse storage (its own formula, include/h2ok_hist.h): 8 bytes per nonzero cell {row, value} instead of 4 bytes per
    cell, and 10 bytes per row (2-byte node + two 4-byte targets)."""
    if sparse_nnz is not None:  # (a forest of `trees` class trees reads the cells once and `trees` row records per row)
        return sparse_nnz * 8.0 + rows * 10.0 * trees + n_nodes * cols * (bins + 1) * 3 * 8.0
    if u8:  # pre-binned codes: one byte per cell
        return rows * cols * 1.0 + rows * 8.0 + n_nodes * cols * (bins + 1) * 3 * 8.0
    return rows * cols * 4.0 + rows * (4.0 + 4.0 + (4.0 if weighted else 0.0)) + n_nodes * cols * (bins + 1) * 3 * 8.0


# dram__bytes_read.sum + dram__bytes_write.sum per histogram launch, from the committed `ncu --set full` captures at
# 100M x 64 (profiles/r01_eng_hist_ncu_full.csv: 26.8-27.5 GB read + 0.01-0.05 GB written; profiles/r01_hist_build_ncu_full.csv:
# 26.5-27.6 GB read + 0.01-0.3 GB written); scaled linearly to the rows of the run.  None for shapes not profiled.
NCU_TRAFFIC_BYTES_PER_CELL = {"engine": 27.13e9 / (100_000_000 * 64), "engine-py": 27.13e9 / (100_000_000 * 64),
                              "stateless": 27.15e9 / (100_000_000 * 64)}


def measured_hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region: NVML from a thread of this process every 5 ms
    (a timed region can be shorter than one nvidia-smi start-up: 32 levels on 8 GPUs are 40 ms), nvidia-smi -lms as the
    fallback when NVML cannot be loaded."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None
        self.nvml, self.handle, self.thread, self.stop_flag = None, None, None, None
        self.sm, self.mx, self.reasons = [], [], set()
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = index
            if vis:  # torch's device index counts the visible devices; NVML counts all of them
                ids = [v.strip() for v in vis.split(",") if v.strip()]
                if index < len(ids) and ids[index].isdigit():
                    phys = int(ids[index])
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.mx_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _sample(self):
        n = self.nvml
        self.sm.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
        self.mx.append(self.mx_mhz)
        try:
            get = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
            r = int(get(self.handle))
        except Exception:
            return
        for name, bit in (("hw_slowdown", 0x8), ("sw_thermal_slowdown", 0x20), ("hw_thermal_slowdown", 0x40),
                          ("sw_power_cap", 0x4)):
            if r & bit:
                self.reasons.add(name)

    def _loop(self):
        while not self.stop_flag.is_set():
            try:
                self._sample()
            except Exception:
                break
            self.stop_flag.wait(0.005)

    def start(self):
        if self.nvml is not None:
            import threading
            self.stop_flag = threading.Event()
            self.thread = threading.Thread(target=self._loop, daemon=True)
            self.thread.start()
            return
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.thread is not None:
            self.stop_flag.set()
            self.thread.join(timeout=2)
            if self.sm:
                out.update(sm_mhz=float(np.median(self.sm)), sm_max_mhz=float(max(self.mx)), samples=len(self.sm),
                           source="nvml, 5 ms period")
            out["reasons"] = sorted(self.reasons)
            return out
        if self.proc is None:
            out["reasons"] = ["nvidia-smi unavailable"]
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1]))
                    mx.append(float(f[2]))
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), samples=len(sm), source="nvidia-smi -lms 100")
        out["reasons"] = sorted(reasons)
        return out


# ------------------------------------------------------------------------------------------------ CPU arms
_REAL_STDOUT = 1


def host_threads():
    """Host cores this process may use.  torch.distributed.run exports OMP_NUM_THREADS=1, so the OpenMP default is not
    the answer under torchrun; the oracle takes the thread count as an argument (num_threads clause)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_sample_rows(wl):
    return int(min(CPU_SAMPLE_ROWS, wl["rows"], max(1000, CPU_SAMPLE_CELLS // wl["cols"])))


def cpu_levels(oracle, wl, sample_rows, threads, max_seconds, levels=None):
    """The oracle's own level loop (fp64, column-parallel like H2O's LocalMR) on the first `sample_rows` rows of the
    workload.  Returns per-level seconds of build_level + find_splits + next_level + apply_splits."""
    from h2ok_hist import abi

    cols, bins = wl["cols"], wl["bins"]
    X, y, _ = oracle.synth(wl["shape"], SEED, 0, sample_rows, cols)
    params = abi.Params.make(nbins=bins, max_depth=wl["depth"])
    descs = oracle.root_descs(X, sample_rows, params).reshape(1, cols)
    nid = np.zeros(sample_rows, np.int32)
    gid, times, nodes = 0, [], []
    t_all = time.perf_counter()
    for depth in range(wl["depth"] if levels is None else levels):
        n = descs.shape[0]
        t0 = time.perf_counter()
        hist, _, min2, maxin = oracle.build_level(X, sample_rows, y, None, nid, descs, bins, threads=threads, want_abs=False)
        splits = oracle.find_splits(hist, descs, params.min_rows, params.min_split_improvement)
        if depth + 1 < wl["depth"]:
            child, ids, nch = oracle.next_level(splits, descs, min2, maxin, params, depth)
        else:
            child, ids, nch = None, np.full(2 * n, -1, np.int32), 0
        oracle.apply_splits(X, sample_rows, splits, ids, gid, nid, None)
        times.append(time.perf_counter() - t0)
        nodes.append(n)
        gid += n
        descs = child
        if nch == 0 or time.perf_counter() - t_all > max_seconds:
            break
    return times, nodes


def run_reference(args, wl):
    """--impl reference: the CPU implementation of the path on the host cores.  The reference tree has no code for
    this path and H2O-3 cannot run here (no JVM, no jar), so this arm is the oracle port with every host thread."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle_lib

    oracle = oracle_lib.load()
    threads = host_threads()
    rows = cpu_sample_rows(wl)
    # every "step" is one level of a depth-`depth` tree over the sample; warm-up levels come first
    need = args.steps + args.warmup
    times = []
    while len(times) < need:
        t, _ = cpu_levels(oracle, wl, rows, threads, max_seconds=600)
        times += t
    timed = times[args.warmup: args.warmup + args.steps]
    sec = float(np.sum(timed))
    value = rows * wl["cols"] * len(timed) / sec
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sec / len(timed), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": wl["name"], "rows": wl["rows"], "cols": wl["cols"], "bins": wl["bins"], "depth": wl["depth"],
                   "sample_rows": rows, "note": "CPU oracle port (H2O-3 itself is not runnable here: no JVM/jar)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"levels of a depth-{wl['depth']} tree on the first {rows} rows x {wl['cols']} cols of the workload"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ GPU arm
class PyLoop:
    """The Python per-step loop (LevelDriver) behind the interface bench uses, for --path stateless / engine-py."""

    def __init__(self, drv, shard):
        self.drv, self.shard = drv, shard

    def begin_tree(self):
        self.drv.prepare_tree()

    def level(self, out=None):
        return self.drv.level()

    def n_nodes_left(self):
        return self.drv.n_nodes_left()

    depth = property(lambda self: self.drv.depth)
    n_nodes = property(lambda self: self.drv.n_nodes)


def cpu_leg(wl, rows_total, gpu_slice_checksum):
    """cpu_baseline: the oracle's level loop on the first CPU_SAMPLE_ROWS rows (a slice, not extrapolated), timed
    single-threaded (north_star's oracle) and with OpenMP over columns on every host core, best of N after one warm-up;
    plus the checksum of the oracle's fixed-point tree on the slice, which must equal the GPU's on the same rows."""
    import oracle_lib
    from h2ok_hist import abi

    oracle = oracle_lib.load()
    threads = host_threads()
    srows = cpu_sample_rows(wl)
    cols, depth = wl["cols"], wl["depth"]

    def best(th, reps, budget):
        vals, t_all = [], time.perf_counter()
        for i in range(reps + 1):
            t, nodes = cpu_levels(oracle, wl, srows, th, max_seconds=budget)
            if i > 0 or reps == 0:
                vals.append(srows * cols * len(t) / float(np.sum(t)))
            if time.perf_counter() - t_all > budget:
                break
        return (max(vals) if vals else srows * cols * len(t) / float(np.sum(t))), len(t)

    v_omp, nlev = best(threads, 5, 12.0)
    v_one, _ = best(1, 1, 15.0)
    X, y, _ = oracle.synth(wl["shape"], SEED, 0, srows, cols)
    levels, _, leaf, _ = oracle_lib.train_tree_q(oracle, X, srows, y, None, abi.Params.make(nbins=wl["bins"], max_depth=depth))
    cks = oracle_lib.tree_checksum(levels, leaf)
    return {"value": v_omp, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{nlev} levels of a depth-{depth} tree on the first {srows} rows x {cols} cols (slice, not "
                      f"extrapolated); OpenMP over columns, best of 5 after one warm-up",
            "single_thread_value": v_one, "single_thread_note": "the same loops on one thread (north_star's oracle), best of 2",
            "checksum_slice_oracle": ["%016x" % c for c in cks],
            "checksum_slice_gpu": None if gpu_slice_checksum is None else ["%016x" % c for c in gpu_slice_checksum],
            "checksum_slice_match": None if gpu_slice_checksum is None else (list(cks) == list(gpu_slice_checksum))}


def run_gpu(args, wl):
    import torch
    import torch.distributed as dist

    from h2ok_hist import abi, capi
    from h2ok_hist.driver import CudaShard, EngineShard, LevelDriver, TreeLoop

    # libraries (NCCL's version banner, for one) may write to stdout: everything but the final line goes to stderr
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N>1 must be launched with torch.distributed.run (one rank per GPU)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    rows_total, cols, bins, depth = wl["rows"], wl["cols"], wl["bins"], wl["depth"]
    row0 = rows_total * rank // world
    rows = rows_total * (rank + 1) // world - row0
    params = abi.Params.make(nbins=bins, max_depth=depth)
    in_library = args.path == "engine"
    K = int(wl.get("k", 1))
    sparse = bool(wl.get("sparse")) and args.path != "stateless"
    nnz = rows * cols * (1.0 - 58982.0 / 65536.0) if sparse else None  # the generator zeroes a cell when hash16 < 58982
    u8 = bool(wl.get("u8")) and args.path != "stateless"
    shard = CudaShard(local) if args.path == "stateless" else EngineShard(local, sparse=sparse, u8=u8)
    if in_library:
        TreeLoop.join_ranks(shard)  # the library's own NCCL communicator (h2ok_hist_comm_init)
    if u8:
        shard.set_data_synth_prebinned(wl["shape"], SEED, row0, rows, cols, nbins=bins)
    else:
        shard.set_data_synth(wl["shape"], SEED, row0, rows, cols)
    # K class trees per boosting iteration: K target vectors.  They are the residuals the first multinomial iteration
    # sees, y_k = 1{label == k} - 1/K with label = the largest of the row's first K columns (what scripts/gbm_c4.py
    # trains on), so that every class tree keeps splitting down to max_depth; independent noise targets (SURVEY.md
    # §8(d)'s y_k = 2 u_k - 1) stop at the root under min_split_improvement and would time nine one-level trees.
    forest = bool(wl.get("forest")) and in_library and K > 1
    ys_all = torch.empty((K, shard.y.numel()), dtype=torch.float32, device=shard.y.device)
    ys_all[0].copy_(shard.y)
    if K > 1:
        best = torch.full((rows,), -float("inf"), device=shard.y.device)
        label = torch.zeros(rows, dtype=torch.int64, device=shard.y.device)
        colbuf = torch.empty(shard.ld, dtype=torch.float32, device=shard.y.device)
        for k in range(K):
            shard.lib.call("h2ok_hist_synth_fill_cols_dev", shard.ctx, wl["shape"], SEED, row0, rows, k, 1, shard.ld, 0,
                           C.c_void_p(colbuf.data_ptr()))
            better = colbuf[:rows] > best
            label[better] = k
            best = torch.maximum(best, colbuf[:rows])
        ys_all.zero_()
        for k in range(K):
            ys_all[k, :rows] = (label == k).float() - 1.0 / K
        del best, label, colbuf
    ys = [ys_all[k] for k in range(K)]
    tree_no = [0]

    def next_targets():
        y = ys[tree_no[0] % K]
        tree_no[0] += 1
        return y
    if in_library:
        # (the checksum's hash kernels run only for the tree built after the timed region, see below)
        loop = TreeLoop(shard, params, checksum=False, time_phases=True, forest=K if forest else 1)
    else:
        drv = LevelDriver(shard, params)
        drv.prepare_data()
        loop = PyLoop(drv, shard)
    barrier()

    # ---- resident path: inputs in HBM when the timed region starts ------------------------------------
    rec = []  # per timed step: (level, n_nodes, hist_ms) — Python-stepped paths only; the library times its own loop

    def run_steps(n_steps, record):
        done = 0
        while done < n_steps:
            if loop.n_nodes_left() == 0:
                begin()
            lvl, n = loop.depth, loop.n_nodes
            loop.level()
            if record and not in_library:
                rec.append((lvl, n, shard.last_hist_ms()))
            done += 1

    def begin():
        if forest:
            loop.begin_forest(ys_all)
        elif in_library:
            loop.begin_tree(y=next_targets())
        else:
            loop.begin_tree()

    begin()
    run_steps(args.warmup, False)
    sampler = ClockSampler(local)
    if in_library:
        loop.reset_phase_stats()
    barrier()
    if rank == 0:
        sampler.start()
    launches0 = shard.launches
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    run_steps(args.steps, True)
    ev1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = float(ms.item())
    launches = shard.launches - launches0
    trees_per_step = K if forest else 1  # a forest level is one level of each of its K class trees
    value = rows_total * cols * trees_per_step * args.steps / (total_ms * 1e-3)
    loop_stats = loop.phase_stats(reset=True) if in_library else None

    # result checksum: a full tree on the timed inputs, built right after the timed region.  Integer sums make every
    # tree on these inputs the same bits, whatever the number of ranks: the lines of a scaling run must agree.
    checksum = None
    if in_library:
        tree_no[0] = 0
        loop.set_options(checksum=True)
        begin()
        while loop.n_nodes_left() > 0:
            loop.level()
        checksum = ["%016x" % c for c in loop.checksum(row0)]
        loop.set_options(checksum=False)

    # roofline of the dominant kernel: algorithmic bytes of THIS rank's launches / their device time (CUDA events
    # recorded inside the library around the histogram kernel of every timed level, read back after the timed region)
    peak, peak_src = measured_hbm_peak()
    level_table, phases = {}, None
    if in_library:
        cnt, nodes, ms = loop_stats
        hist_ms = float(ms[:, 0].sum())
        abytes = float(sum(int(cnt[d]) * algo_bytes_level(rows, cols, int(nodes[d]), bins, sparse_nnz=nnz, u8=u8, trees=trees_per_step) for d in range(len(cnt))))
        for d in range(len(cnt)):
            if cnt[d]:
                t = float(ms[d, 0] / cnt[d])
                level_table[str(d)] = {
                    "nodes": int(nodes[d]), "hist_ms": round(t, 4),
                    "frac": round(algo_bytes_level(rows, cols, int(nodes[d]), bins, sparse_nnz=nnz, u8=u8, trees=trees_per_step) / (t * 1e-3) / 1e9 / peak, 4),
                    "reduce_ms": round(float(ms[d, 2] / cnt[d]), 4), "search_ms": round(float(ms[d, 3] / cnt[d]), 4),
                    "route_ms": round(float(ms[d, 4] / cnt[d]), 4)}
        tot = ms.sum(axis=0) / max(1, int(cnt.sum()))
        step_ms = total_ms / args.steps
        phases = {"hist_kernel_ms": round(float(tot[0]), 4), "clear_and_hist_ms": round(float(tot[1]), 4),
                  "reduce_ms": round(float(tot[2]), 4), "search_children_ms": round(float(tot[3]), 4),
                  "route_regroup_ms": round(float(tot[4]), 4), "other_ms": round(float(step_ms - tot[1:].sum()), 4),
                  "levels_timed": int(cnt.sum()),
                  "note": "device time per level, CUDA events inside the library around the phases of every level of the "
                          "timed region (this rank), read back afterwards; other = per-tree set-up (absmax, quantise, root "
                          "records) and gaps"}
    else:
        hist_ms = float(sum(r[2] for r in rec))
        abytes = float(sum(algo_bytes_level(rows, cols, r[1], bins) for r in rec))
        per_level = {}
        for lvl, n, t in rec:
            per_level.setdefault(lvl, []).append((n, t))
        for k, v in sorted(per_level.items()):
            t = float(np.mean([x[1] for x in v]))
            level_table[str(k)] = {"nodes": v[0][0], "hist_ms": round(t, 4),
                                   "frac": round(algo_bytes_level(rows, cols, v[0][0], bins) / (t * 1e-3) / 1e9 / peak, 4)}
    achieved = abytes / (hist_ms * 1e-3) / 1e9 if hist_ms > 0 else 0.0
    whole_level_frac = abytes / (total_ms * 1e-3) / 1e9 / peak

    # ---- e2e: host buffers in, host buffers out, copies inside the timed region ---------------------------
    e2e = None
    if not args.no_e2e and in_library and not forest:
        e2e = run_e2e_loop(args, wl, shard, loop, world, rank, rows_total, rows, barrier)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        gpu_slice = None
        if in_library:
            srows = cpu_sample_rows(wl)
            s2 = EngineShard(local, sparse=sparse)
            s2.set_data_synth(wl["shape"], SEED, 0, srows, cols)  # (u8 workloads: the slice check runs on the fp32 matrix)
            l2 = TreeLoop(s2, params, checksum=True)
            l2.train_tree()
            gpu_slice = l2.checksum(0)
            s2.close()
        cpu = cpu_leg(wl, rows_total, gpu_slice)

    if rank == 0:
        sys.stdout.flush()
        os.dup2(_REAL_STDOUT, 1)  # the JSON line is the only thing this process writes to its real stdout
        kern = "hist_build_kernel" if args.path == "stateless" else "eng_hist_kernel"
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64 binning, i32/i64 fixed-point sums", "data": "synthetic",
            "config": {"workload": wl["name"], "rows": rows_total, "cols": cols, "bins": bins, "depth": depth,
                       "rows_per_gpu": rows, "path": args.path,
                       "step": "one tree level (levels cycle 0..depth-1 of real trees); " +
                               ("one h2ok_hist_loop_level call per step, reduce over ranks by the library's NCCL calls"
                                if in_library else "Python per-step loop"),
                       "l2": "inputs larger than L2 (%.1f GB per GPU per level)" % (rows * cols * 4 / 1e9),
                       "per_level": level_table, "phases_ms": phases, "whole_level_frac": round(whole_level_frac, 4),
                       **({"k_classes": K, "storage": "sparse (H2OK_TREE_SPARSE)" if sparse else "dense fp32",
                           "forest": forest,
                           "fused_level_ms": round((1 if forest else K) * total_ms / args.steps, 3),
                           "fused_level_note": "one level of all K class trees = K steps; SURVEY.md §8(d)'s floor for it is one "
                                               "read of the DENSE matrix (80.8 GB): frac_of_dense_fused_floor compares with that",
                           "frac_of_dense_fused_floor": round((rows * cols * 4.0 + K * rows * 8.0) /
                                                              ((1 if forest else K) * total_ms / args.steps * 1e-3) / 1e9 / peak, 4)}
                          if K > 1 else {})},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": (NCU_TRAFFIC_BYTES_PER_CELL[args.path] * rows * cols / 1e9
                                     if args.workload == "c3" and args.path in NCU_TRAFFIC_BYTES_PER_CELL else None),
                         "traffic_unit": "GB per launch: ncu dram read+write of the N=1 capture at 100M rows "
                                         "(profiles/, see README), scaled linearly to this rank's rows",
                         "kernel": kern, "peak_source": peak_src,
                         "note": "algorithmic bytes of the timed levels / CUDA-event time of their histogram launches"},
            "checksum": checksum, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    shard.close()
    if world > 1:
        dist.destroy_process_group()


def run_e2e_loop(args, wl, shard, loop, world, rank, rows_total, rows, barrier):
    """The same metric through the host-buffer side of the C ABI (what an H2O-side binding calls): per tree the targets
    y come from host memory (h2ok_hist_loop_begin_tree_host: H2D inside the timed region), per level
    h2ok_hist_loop_level fills host buffers with the histogram, observed ranges, splits, child descriptors and child
    ids (D2H inside the timed region; with several ranks every rank receives the (node, column) pairs it reduced, so
    the reduced histogram reaches host memory exactly once).  X stays resident between levels and trees, as it does in a
    training job; its one-time upload through h2ok_hist_set_data_host is measured beside it on a bounded sample."""
    import torch

    from h2ok_hist import abi, capi

    cols, bins, depth = wl["cols"], wl["bins"], wl["depth"]
    nb1 = bins + 1
    max_nodes = 2 ** (depth - 1)
    yh = shard.y[:rows].cpu().pin_memory()  # (K > 1: every class tree re-uploads a target vector of the same size)

    def pinned(shape, dtype):
        return torch.empty(shape, dtype=dtype).pin_memory()

    bufs = dict(hist=pinned((max_nodes, cols, nb1, 3), torch.float64), min2=pinned((max_nodes, cols), torch.float32),
                maxin=pinned((max_nodes, cols), torch.float32), splits=pinned((max_nodes, abi.SPLIT.itemsize), torch.uint8),
                child_descs=pinned((max_nodes, cols, abi.COL_DESC.itemsize), torch.uint8),
                child_ids=pinned((2 * max_nodes,), torch.int32))
    out = abi.LevelOut(*[C.c_void_p(bufs[k].data_ptr()) for k in ("hist", "min2", "maxin", "splits", "child_descs", "child_ids")])
    st = dict(h2d=0, d2h=0)

    def steps(k):
        done = 0
        while done < k:
            if loop.n_nodes_left() == 0:
                loop.begin_tree_host(yh)
                st["h2d"] += rows * 4
            n = loop.n_nodes
            _, nch = loop.level(out)
            pairs = n * cols
            share = -(-pairs // world)  # the pairs this rank reduced
            st["d2h"] += min(share, pairs) * nb1 * 24 + pairs * 8 + n * abi.SPLIT.itemsize + 2 * n * 4 + 4 + \
                nch * cols * abi.COL_DESC.itemsize
            done += 1

    steps(min(args.warmup, depth))
    while loop.n_nodes_left() > 0:  # the timed region starts at a tree boundary so that the y upload is counted
        loop.level(out)
    st.update(h2d=0, d2h=0)
    barrier()
    t0 = time.perf_counter()
    steps(args.steps)
    torch.cuda.synchronize()
    sec = time.perf_counter() - t0  # loop_level is complete on return as far as its host buffers go
    barrier()
    t = torch.tensor([sec], dtype=torch.float64, device="cuda")
    if world > 1:
        import torch.distributed as dist

        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    sec = float(t.item())
    up = None
    if rank == 0 and world == 1:
        up = measure_upload(wl, int(min(rows, max(1000, 640_000_000 // cols))))
    return {"value": rows_total * cols * args.steps / sec, "unit": UNIT, "h2d_bytes_per_step": st["h2d"] / args.steps,
            "d2h_bytes_per_step": st["d2h"] / args.steps, "ms_per_step": sec * 1e3 / args.steps,
            "path": "C ABI host-buffer calls: h2ok_hist_loop_begin_tree_host per tree (y from host memory), "
                    "h2ok_hist_loop_level per level with host buffers for histogram, ranges, splits, child descriptors, child ids; "
                    "host wall clock, max over ranks",
            "x_upload": up}


def measure_upload(wl, srows):
    """One-time h2ok_hist_set_data_host of a bounded sample of the workload from pageable host memory."""
    import torch

    from h2ok_hist import capi

    lib = capi.load()
    cols = wl["cols"]
    Xh = torch.empty((cols, srows), dtype=torch.float32)
    Xh.uniform_(0, 1)
    h = C.c_void_p()
    lib.call("h2ok_hist_create", 0, C.byref(h))
    colptrs = (C.c_void_p * cols)(*[Xh[c].data_ptr() for c in range(cols)])
    t0 = time.perf_counter()
    lib.call("h2ok_hist_set_data_host", h, colptrs, srows, cols)
    sec = time.perf_counter() - t0
    lib.call("h2ok_hist_destroy", h)
    return {"note": "one-time h2ok_hist_set_data_host (pageable host columns -> row-major resident copy + routing codes), "
                    "outside the timed region; measured on a sample",
            "sample_rows": srows, "bytes": srows * cols * 4, "ms": sec * 1e3, "gbs": srows * cols * 4 / sec / 1e9}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0, help="override the workload's row count (debug only)")
    ap.add_argument("--path", default="engine", choices=["engine", "engine-py", "stateless"],
                    help="engine (default): resident row-wise tree engine driven by the level loop inside the library "
                         "(h2ok_hist_loop_level); engine-py: the same kernels stepped from Python (LevelDriver); "
                         "stateless: per-call kernels on column-major rows in place")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (profiling runs only)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    wl = dict(WORKLOADS[args.workload])
    if args.rows:
        wl["rows"] = args.rows
        wl["name"] += f" [rows overridden to {args.rows}]"
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_gpu(args, wl)


if __name__ == "__main__":
    main()
