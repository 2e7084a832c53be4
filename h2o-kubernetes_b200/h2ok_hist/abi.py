"""ctypes / numpy mirrors of the structs in include/h2ok_hist.h (one definition for product and tests).

The reference has no FFI; these layouts are defined by include/h2ok_hist.h and must match it
byte for byte — tests/test_abi.py checks sizes and offsets against the compiled library.
"""
import ctypes as C

import numpy as np

OK, ERR_ARG, ERR_CUDA, ERR_STATE, ERR_UNSUPPORTED = 0, 1, 2, 3, 4
MAX_BINS = 1024
NA_NONE, NA_VS_REST, NA_LEFT, NA_RIGHT = 0, 1, 2, 3
SHAPE_HIGGS28, SHAPE_WIDE64, SHAPE_SPARSE = 2, 3, 4
DIST_GAUSSIAN, DIST_BERNOULLI, DIST_MULTINOMIAL = 0, 1, 2
REDUCE_SCATTER, REDUCE_ALL = 0, 1
TREE_WEIGHTED, TREE_SPARSE, TREE_U8 = 1, 2, 4
COMM_ID_BYTES = 128

COL_DESC = np.dtype(
    [("min", "<f8"), ("maxex", "<f8"), ("step", "<f8"), ("nbin", "<i4"), ("is_int", "<i4")], align=True
)
SPLIT = np.dtype(
    [
        ("col", "<i4"), ("bin", "<i4"), ("na_dir", "<i4"), ("splat", "<f4"),
        ("se_before", "<f8"), ("se_left", "<f8"), ("se_right", "<f8"),
        ("n_left", "<f8"), ("n_right", "<f8"), ("pred_left", "<f8"), ("pred_right", "<f8"),
        ("equal", "<i4"), ("pad", "<i4"), ("cat_bits", "<u4", (32,)),
    ],
    align=True,
)
CAT_WORDS = 32
assert COL_DESC.itemsize == 32 and SPLIT.itemsize == 208


class Quant(C.Structure):
    _fields_ = [("scale_w", C.c_double), ("scale_wy", C.c_double), ("scale_wyy", C.c_double),
                ("weighted", C.c_int32), ("qbits", C.c_int32)]


class Params(C.Structure):
    _fields_ = [("nbins", C.c_int32), ("nbins_top_level", C.c_int32), ("max_depth", C.c_int32), ("pad", C.c_int32),
                ("min_rows", C.c_double), ("min_split_improvement", C.c_double)]

    @classmethod
    def make(cls, nbins=256, nbins_top_level=None, max_depth=6, min_rows=10.0, min_split_improvement=1e-5):
        return cls(nbins, nbins if nbins_top_level is None else nbins_top_level, max_depth, 0,
                   float(min_rows), float(min_split_improvement))


class Stats(C.Structure):
    _fields_ = [("build_ms", C.c_float), ("launches", C.c_int32), ("passes", C.c_int32), ("copies", C.c_int32),
                ("range_errors", C.c_int32), ("pad", C.c_int32), ("algo_bytes", C.c_double)]


class LevelOut(C.Structure):
    """h2ok_level_out: host buffers h2ok_hist_loop_level fills (NULL = not wanted)."""
    _fields_ = [("hist", C.c_void_p), ("min2", C.c_void_p), ("maxin", C.c_void_p), ("splits", C.c_void_p),
                ("child_descs", C.c_void_p), ("child_ids", C.c_void_p)]


class LoopView(C.Structure):
    """h2ok_loop_view: where the loop stands + device pointers to the level that ran last."""
    _fields_ = [("depth", C.c_int32), ("n_nodes", C.c_int32), ("gid_base", C.c_int32),
                ("max_nodes", C.c_int32), ("max_bins", C.c_int32), ("cols", C.c_int32), ("rank", C.c_int32),
                ("n_ranks", C.c_int32), ("reduce_mode", C.c_int32),
                ("done_depth", C.c_int32), ("done_nodes", C.c_int32), ("done_children", C.c_int32),
                ("done_gid_base", C.c_int32), ("n_trees", C.c_int32),
                ("done_pair0", C.c_int64), ("done_pair1", C.c_int64), ("quant", Quant),
                ("d_descs", C.c_void_p), ("d_wq", C.c_void_p), ("d_wyq", C.c_void_p), ("d_wyyq", C.c_void_p),
                ("d_leaf", C.c_void_p), ("d_done_descs", C.c_void_p), ("d_done_acc", C.c_void_p),
                ("d_done_min2_enc", C.c_void_p), ("d_done_maxin_enc", C.c_void_p), ("d_done_splits", C.c_void_p),
                ("d_done_child_ids", C.c_void_p)]


def ptr(a):
    """Raw address of a numpy array (or None → NULL)."""
    return None if a is None else C.c_void_p(a.ctypes.data)


def decode_f32(enc):
    """Inverse of the order-preserving uint32 encoding used for observed min/max (see csrc/common.cuh)."""
    enc = np.asarray(enc, dtype=np.uint32)
    bits = np.where(enc & np.uint32(0x80000000), enc & np.uint32(0x7FFFFFFF), ~enc)
    return bits.astype(np.uint32).view(np.float32)


def encode_f32(x):
    bits = np.asarray(x, dtype=np.float32).view(np.uint32)
    return np.where(bits & np.uint32(0x80000000), ~bits, bits | np.uint32(0x80000000)).astype(np.uint32)
