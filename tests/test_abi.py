"""CPU tests of the boundary: the library loads without a GPU, exports every symbol include/h2ok_hist.h declares,
its struct layouts match the Python mirrors, and it fails loudly (status + message) when no device is present."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from h2ok_hist import abi, capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "h2ok_hist.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(h2ok_[a-z0-9_]+)\s*\(", src)))


def test_library_loads_and_exports_every_declared_symbol():
    if not os.path.exists(capi.LIB_PATH):
        import __graft_entry__

        __graft_entry__.build()
    lib = capi.load()
    names = declared_functions()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib.cdll, n), f"{n} declared in h2ok_hist.h but not exported"
    bound = set(capi.SIGNATURES) | {"h2ok_last_error"}
    assert set(names) == bound, set(names) ^ bound
    assert lib.cdll.h2ok_hist_abi_version() == 2


def test_struct_layouts_match_header(tmp_path):
    prog = tmp_path / "layout.c"
    prog.write_text(
        '#include <stdio.h>\n#include <stddef.h>\n#include "h2ok_hist.h"\n'
        "int main(void){\n"
        'printf("%zu %zu %zu %zu %zu\\n", sizeof(h2ok_col_desc), sizeof(h2ok_split), sizeof(h2ok_quant), sizeof(h2ok_params), sizeof(h2ok_stats));\n'
        'printf("%zu %zu %zu %zu\\n", offsetof(h2ok_col_desc, step), offsetof(h2ok_col_desc, nbin), offsetof(h2ok_split, splat), offsetof(h2ok_split, pred_right));\n'
        'printf("%zu %zu %zu\\n", offsetof(h2ok_quant, weighted), offsetof(h2ok_params, min_rows), offsetof(h2ok_stats, algo_bytes));\n'
        "return 0;}\n"
    )
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(prog), "-o", str(exe)])
    out = subprocess.check_output([str(exe)]).decode().split()
    sizes = [int(x) for x in out]
    assert sizes[:5] == [abi.COL_DESC.itemsize, abi.SPLIT.itemsize, C.sizeof(abi.Quant), C.sizeof(abi.Params), C.sizeof(abi.Stats)]
    assert sizes[5:9] == [abi.COL_DESC.fields["step"][1], abi.COL_DESC.fields["nbin"][1], abi.SPLIT.fields["splat"][1],
                          abi.SPLIT.fields["pred_right"][1]]
    assert sizes[9:] == [abi.Quant.weighted.offset, abi.Params.min_rows.offset, abi.Stats.algo_bytes.offset]


def test_encoding_roundtrip_and_order():
    x = np.array([-np.inf, -3.5, -0.0, 0.0, 1e-30, 2.0, np.inf], np.float32)
    e = abi.encode_f32(x)
    assert (np.diff(e.astype(np.int64)) > 0).all()
    assert np.array_equal(abi.decode_f32(e).view(np.uint32), x.view(np.uint32))


def test_host_helpers_need_no_gpu(oracle):
    lib = capi.load()
    p = abi.Params.make(nbins=16)
    lo = np.array([-1.0, 5.0, 0.0, np.inf], np.float32)
    hi = np.array([1.0, 5.0, 2.0, -np.inf], np.float32)
    ii = np.array([0, 0, 1, 0], np.int32)
    out = np.zeros(4, abi.COL_DESC)
    lib.call("h2ok_hist_root_descs_from_ranges", abi.ptr(lo), abi.ptr(hi), 4, C.byref(p), abi.ptr(ii), abi.ptr(out))
    assert out["nbin"].tolist() == [16, 0, 3, 0]
    X = np.zeros((4, 8), np.float32)
    X[0, :2] = [-1, 1]
    X[1] = 5
    X[2, :3] = [0, 1, 2]
    X[3] = np.nan
    X[0, 2:] = 0.5
    want = oracle.root_descs(X, 8, p, ii)
    assert out.tobytes() == want.tobytes()  # product's host arithmetic == oracle's, byte for byte
    m = np.array([1.5, 0.7, 3.0])
    q = abi.Quant()
    lib.call("h2ok_hist_choose_quant", abi.ptr(m), 1, 25, C.byref(q))
    qo = abi.Quant()
    oracle.lib.h2ok_oracle_choose_quant(abi.ptr(m), 1, 25, C.byref(qo))
    assert (q.scale_w, q.scale_wy, q.scale_wyy) == (qo.scale_w, qo.scale_wy, qo.scale_wyy) == (2.0**24, 2.0**25, 2.0**23)
    # a categorical column (is_int == 2) is binned like an integer column; its codes must fit a split's bit set
    ii[2] = 2
    lib.call("h2ok_hist_root_descs_from_ranges", abi.ptr(lo), abi.ptr(hi), 4, C.byref(p), abi.ptr(ii), abi.ptr(out))
    assert out["nbin"].tolist() == [16, 0, 3, 0] and out["is_int"].tolist() == [0, 0, 2, 0]
    assert out.tobytes() == oracle.root_descs(X, 8, p, ii).tobytes()
    hi2 = hi.copy()
    hi2[2] = 1024.0
    with pytest.raises(capi.H2okError) as e:
        lib.call("h2ok_hist_root_descs_from_ranges", abi.ptr(lo), abi.ptr(hi2), 4, C.byref(p), abi.ptr(ii), abi.ptr(out))
    assert e.value.code == abi.ERR_UNSUPPORTED and "categorical" in str(e.value)
    ii[2] = 3
    with pytest.raises(capi.H2okError) as e:
        lib.call("h2ok_hist_root_descs_from_ranges", abi.ptr(lo), abi.ptr(hi), 4, C.byref(p), abi.ptr(ii), abi.ptr(out))
    assert e.value.code == abi.ERR_ARG


def test_fails_loudly_without_a_device():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    lib = capi.load()
    h = C.c_void_p()
    with pytest.raises(capi.H2okError) as e:
        lib.call("h2ok_hist_create", 0, C.byref(h))
    assert e.value.code == abi.ERR_CUDA  # no silent CPU path


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "h2o-kubernetes_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle_lib" not in txt and "liboracle" not in txt and "hist_oracle.h" not in txt, f
