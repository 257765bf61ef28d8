"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/mxmoe_b200.h declares, and the ctypes mirror agrees with the header (no compute calls)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
HEADER = os.path.join(ROOT, "include", "mxmoe_b200.h")
LIB = os.path.join(ROOT, "mxmoe_b200", "lib", "libmxmoe_b200.so")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mxb_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(LIB):
        import __graft_entry__ as g
        g.build()
    return ctypes.CDLL(LIB)


def test_header_declares_expected_surface():
    syms = declared_symbols()
    for s in ("mxb_create", "mxb_group_gemm", "mxb_group_gemm_reflayout", "mxb_permute", "mxb_permute_static", "mxb_quant_act",
              "mxb_silu_mul_quant", "mxb_combine", "mxb_quant_pack_weight", "mxb_repack_from_ref", "mxb_select_tiles"):
        assert s in syms


def test_library_exports_every_declared_symbol(lib):
    for s in declared_symbols():
        assert hasattr(lib, s), f"{s} declared in the header but not exported"


def test_ctypes_mirror_matches_header(lib):
    from mxmoe_b200 import ops
    names = sorted(n for n, _, _ in ops.EXPORTS)
    assert names == declared_symbols()
    assert ctypes.sizeof(ops.MxbProblem) == 6 * 8 + 2 * 8 + 9 * 4 + 4 + 16 + 8  # 6 pointers, 2 int64, 9 int32, padding, c_row, a_rows, lds_a
    assert ctypes.sizeof(ops.MxbRowDest) == 32
    assert ctypes.sizeof(ops.MxbActQ) == 8
    assert ctypes.sizeof(ops.MxbTileChoice) == 24
    # ... and against the structs as the library was compiled
    for which, st in enumerate((ops.MxbProblem, ops.MxbActQ, ops.MxbRowDest, ops.MxbTileChoice)):
        assert lib.mxb_struct_size(which) == ctypes.sizeof(st)


def test_pure_host_entry_points(lib):
    lib.mxb_packed_bytes.restype = ctypes.c_int64
    lib.mxb_scale_elems.restype = ctypes.c_int64
    assert lib.mxb_packed_bytes(2816, 2048, 4) == 2816 * 2048 // 2
    assert lib.mxb_scale_elems(2816, 2048, 4, 16, 128, 0) == 16 * 2816 * 2
    assert lib.mxb_scale_elems(2816, 2048, 8, 8, -1, 1) == 2816
    lib.mxb_version.restype = ctypes.c_char_p
    assert b"sm_100a" in lib.mxb_version()


def test_no_cpu_fallback_without_gpu():
    """On a box without a GPU the product must fail loudly, not fall back."""
    import torch
    from mxmoe_b200 import ops
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(ops.MxbError):
        ops.Handle(0)
