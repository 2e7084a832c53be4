"""ctypes binding of libh2ok_hist.so — exactly the entry points declared in include/h2ok_hist.h.

There is no fallback: if the CUDA library is missing or fails to load, importing callers get an exception.
"""
import ctypes as C
import os

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libh2ok_hist.so")

i32, i64, u64, f64, vp = C.c_int32, C.c_int64, C.c_uint64, C.c_double, C.c_void_p
pQ, pP, pS = C.POINTER(abi.Quant), C.POINTER(abi.Params), C.POINTER(abi.Stats)
pLO, pLV = C.POINTER(abi.LevelOut), C.POINTER(abi.LoopView)

# name -> argtypes; every function returns int32 status except h2ok_last_error
SIGNATURES = {
    "h2ok_hist_create": [i32, C.POINTER(vp)],
    "h2ok_hist_destroy": [vp],
    "h2ok_hist_abi_version": [],
    "h2ok_hist_set_stream": [vp, vp, i32],
    "h2ok_hist_sync": [vp],
    "h2ok_hist_copy_to_host": [vp, vp, vp, C.c_size_t],
    "h2ok_hist_synth_fill_dev": [vp, i32, u64, i64, i64, i32, i64, i32, i32, vp, vp, vp],
    "h2ok_hist_synth_fill_cols_dev": [vp, i32, u64, i64, i64, i32, i32, i64, i32, vp],
    "h2ok_hist_col_ranges_dev": [vp, vp, i64, i64, i32, vp, vp],
    "h2ok_hist_target_absmax_dev": [vp, vp, vp, i64, vp],
    "h2ok_hist_quantize_rows_dev": [vp, vp, vp, i64, pQ, vp, vp, vp],
    "h2ok_hist_bin_dev": [vp, vp, i64, vp, vp],
    "h2ok_hist_build_level_dev": [vp, vp, i64, i64, i32, vp, vp, vp, vp, i32, vp, i32, i32, vp, vp, vp],
    "h2ok_hist_build_level_passes": [i32, i32, i64, i32],
    "h2ok_hist_finalize_level_dev": [vp, vp, vp, vp, i32, i32, i32, pQ, vp, vp, vp],
    "h2ok_hist_find_splits_dev": [vp, vp, vp, i32, i32, i32, f64, f64, vp],
    "h2ok_hist_next_level_dev": [vp, vp, vp, vp, vp, i32, i32, pP, i32, vp, vp, vp],
    "h2ok_hist_next_level_noalloc_dev": [vp, vp, vp, vp, vp, i32, i32, pP, i32, vp, vp, vp],
    "h2ok_hist_apply_splits_dev": [vp, vp, i64, i64, vp, vp, i32, i32, vp, vp],
    "h2ok_hist_tree_create": [vp, i64, i32, i32],
    "h2ok_hist_tree_load_cols_dev": [vp, i32, i32, vp, i64],
    "h2ok_hist_tree_destroy": [vp],
    "h2ok_hist_tree_load_bins_dev": [vp, i32, i32, vp, i64],
    "h2ok_hist_prebin_dev": [vp, vp, i64, i64, i32, vp, vp, i64],
    "h2ok_hist_tree_begin_dev": [vp, vp, vp, vp, vp],
    "h2ok_hist_tree_build_level_dev": [vp, i32, vp, i32, i32, vp, vp, vp],
    "h2ok_hist_tree_level_bins": [vp, i32],
    "h2ok_hist_tree_apply_splits_dev": [vp, vp, vp, i32, i32, i32],
    "h2ok_hist_tree_nid_dev": [vp, vp],
    "h2ok_hist_tree_leaf_dev": [vp, vp],
    "h2ok_hist_tree_load_synth": [vp, i32, u64, i64, i32],
    "h2ok_hist_comm_unique_id": [vp],
    "h2ok_hist_comm_init": [vp, vp, i32, i32],
    "h2ok_hist_loop_create": [vp, pP, vp],
    "h2ok_hist_loop_create_forest": [vp, pP, vp, i32],
    "h2ok_hist_loop_begin_forest_dev": [vp, vp, i64],
    "h2ok_hist_loop_forest_quants": [vp, vp],
    "h2ok_hist_loop_destroy": [vp],
    "h2ok_hist_loop_root_descs": [vp, vp],
    "h2ok_hist_loop_set_reduce": [vp, i32],
    "h2ok_hist_loop_set_options": [vp, i32, i32],
    "h2ok_hist_loop_begin_tree_dev": [vp, vp, vp],
    "h2ok_hist_loop_begin_tree_host": [vp, vp, vp],
    "h2ok_hist_loop_begin_tree_synth": [vp, u64, i64, i32],
    "h2ok_hist_loop_level": [vp, pLO, C.POINTER(i32), C.POINTER(i32)],
    "h2ok_hist_loop_view": [vp, pLV],
    "h2ok_hist_loop_phase_stats": [vp, i32, i32, vp, vp, vp],
    "h2ok_hist_loop_checksum": [vp, i64, vp],
    "h2ok_hist_group_create": [vp, i32, C.POINTER(vp)],
    "h2ok_hist_group_destroy": [vp],
    "h2ok_hist_group_size": [vp],
    "h2ok_hist_group_member": [vp, i32, C.POINTER(vp), C.POINTER(i64), C.POINTER(i64)],
    "h2ok_hist_group_load_synth": [vp, i32, u64, i64, i32, i32],
    "h2ok_hist_group_set_data_host": [vp, C.POINTER(vp), i64, i32],
    "h2ok_hist_group_loop_create": [vp, pP, vp],
    "h2ok_hist_group_set_options": [vp, i32, i32, i32],
    "h2ok_hist_group_begin_tree_synth": [vp, u64, i32],
    "h2ok_hist_group_begin_tree_host": [vp, vp, vp],
    "h2ok_hist_group_level": [vp, pLO, C.POINTER(i32), C.POINTER(i32)],
    "h2ok_hist_group_checksum": [vp, vp],
    "h2ok_hist_group_hist_ms": [vp, vp],
    "h2ok_hist_group_sync": [vp],
    "h2ok_hist_gbm_residual_dev": [vp, i32, vp, vp, vp, i64, f64, vp, vp],
    "h2ok_hist_gbm_residual_multi_dev": [vp, i32, vp, vp, i64, vp, i64, f64, vp, vp],
    "h2ok_hist_gbm_gamma_pass_dev": [vp, vp, vp, vp, i64, i32, vp, vp],
    "h2ok_hist_gbm_add_gamma_dev": [vp, vp, i64, vp, i32, f64, vp],
    "h2ok_hist_root_descs_from_ranges": [vp, vp, i32, pP, vp, vp],
    "h2ok_hist_choose_quant": [vp, i32, i32, pQ],
    "h2ok_hist_set_data_host": [vp, C.POINTER(vp), i64, i32],
    "h2ok_hist_set_targets_host": [vp, vp, vp],
    "h2ok_hist_root_descs": [vp, pP, vp, vp],
    "h2ok_hist_build_level": [vp, i32, vp, i32, vp, vp, vp],
    "h2ok_hist_find_splits": [vp, f64, f64, vp],
    "h2ok_hist_next_level": [vp, pP, i32, vp, vp, C.POINTER(i32)],
    "h2ok_hist_apply_splits": [vp],
    "h2ok_hist_get_nid_host": [vp, vp, vp],
    "h2ok_hist_get_quant": [vp, pQ],
    "h2ok_hist_last_level_stats": [vp, pS],
}


class H2okError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"h2ok_hist status {code}: {msg}")
        self.code = code


class Lib:
    def __init__(self, cdll):
        self.cdll = cdll
        cdll.h2ok_last_error.restype = C.c_char_p
        cdll.h2ok_last_error.argtypes = []
        for name, args in SIGNATURES.items():
            fn = getattr(cdll, name)  # AttributeError if the library does not export a declared symbol
            fn.restype = i32
            fn.argtypes = args

    def value(self, name, *args):
        """For the few entry points that return a number instead of a status."""
        return getattr(self.cdll, name)(*args)

    def call(self, name, *args):
        rc = getattr(self.cdll, name)(*args)
        if rc != abi.OK:
            raise H2okError(rc, self.cdll.h2ok_last_error().decode())
        return rc

    def last_error(self):
        return self.cdll.h2ok_last_error().decode()


_LIB = None


def load():
    """Load the CUDA library or raise — the product path has no CPU implementation to fall back to."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `make -C h2o-kubernetes_b200/csrc` "
                "(or __graft_entry__.build()); h2ok_hist has no CPU fallback"
            )
        _LIB = Lib(C.CDLL(LIB_PATH))
    return _LIB
