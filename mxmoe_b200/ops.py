"""ctypes binding of libmxmoe_b200.so (include/mxmoe_b200.h) and the host-side mirror of the
reference's operator module.

The reference's (dead) pybind module `mxmoe_ops` (mxmoe/kernels/src/ref_bind.cu:759-766) exposes
`quant_inp_act`, `gg_mxmoe_share_fused` and `silu_mul_then_quant` with
`qparams_per_exp: list[tuple(a_bits, w_bits, gsize, sym)]`; the functions of the same names below keep
that argument meaning.  PyTorch is used only for device memory and streams: every call passes
`tensor.data_ptr()` and the current CUDA stream to the C-ABI.  There is NO CPU fallback: if the
shared library is missing or the device is not sm_100 the import of a symbol / creation of the
handle raises.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import torch

# MXB_LIB: load another build of the same library (A/B runs of a kernel variant); never a different implementation
_LIB_PATH = os.environ.get("MXB_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libmxmoe_b200.so")


EPI_SILU_MUL = 1  # MXB_EPI_SILU_MUL
PRO_SILU_MUL = 2  # MXB_PRO_SILU_MUL
BITS_E4M3 = -8    # MXB_BITS_E4M3: a_bits / w_bits code of 8-bit E4M3 containers


class MxbError(RuntimeError):
    pass


class MxbProblem(C.Structure):
    _fields_ = [
        ("A", C.c_void_p), ("B", C.c_void_p), ("scale_a", C.c_void_p), ("scale_b", C.c_void_p),
        ("C", C.c_void_p), ("m_range", C.c_void_p),
        ("lda", C.c_int64), ("ldc", C.c_int64),
        ("M", C.c_int32), ("N", C.c_int32), ("K", C.c_int32),
        ("a_bits", C.c_int32), ("w_bits", C.c_int32), ("gsize", C.c_int32), ("sym", C.c_int32),
        ("tile_n", C.c_int32), ("flags", C.c_int32),
        ("c_row", C.c_void_p), ("a_rows", C.c_void_p),
        ("lds_a", C.c_int64),
    ]


class MxbActQ(C.Structure):
    _fields_ = [("a_bits", C.c_int32), ("gsize", C.c_int32)]


class MxbRowDest(C.Structure):
    _fields_ = [("q8", C.c_void_p), ("f16", C.c_void_p), ("scales", C.c_void_p), ("row0", C.c_int64)]


class MxbTileChoice(C.Structure):
    _fields_ = [("tile_m", C.c_int32), ("tile_n", C.c_int32), ("tile_k", C.c_int32), ("stages", C.c_int32),
                ("path", C.c_int32), ("est_us", C.c_float)]


# every symbol include/mxmoe_b200.h declares: (name, restype, argtypes)
_P, _I, _I64 = C.c_void_p, C.c_int, C.c_int64
EXPORTS = [
    ("mxb_create", _I, [_I, C.POINTER(_P)]),
    ("mxb_destroy", _I, [_P]),
    ("mxb_last_error", C.c_char_p, []),
    ("mxb_sm_count", _I, [_P]),
    ("mxb_version", C.c_char_p, []),
    ("mxb_struct_size", _I, [_I]),
    ("mxb_set_gemm_path", _I, [_P, _I]),
    ("mxb_set_sm_limit", _I, [_P, _I]),
    ("mxb_group_gemm", _I, [_P, C.POINTER(MxbProblem), _I, _P]),
    ("mxb_group_gemm_debug_acc", _I, [_P, C.POINTER(MxbProblem), _I, C.POINTER(_P), _P]),
    ("mxb_group_gemm_reflayout", _I, [_P] + [_P] * 14 + [_I, _P]),
    ("mxb_set_ref_operands", _I, [_P, _I]),
    ("mxb_repack_act_from_ref", _I, [_P, _P, _I64, _I, _I, _P, _P]),
    ("mxb_permute", _I, [_P, _P, _I, _I, _I, _P, _P, _P, _P]),
    ("mxb_permute_static", _I, [_I, _P, _I, _I, _I, _P, _P, _P, _P]),
    ("mxb_quant_act", _I, [_P, _P, _I, _P, _I, _P, C.POINTER(MxbActQ), _I, _I, _P, _P, _P, _I, _P]),
    ("mxb_silu_mul_quant", _I, [_P, _P, _I, _I, _P, C.POINTER(MxbActQ), _I, _I, _P, _P, _P, _I, _P]),
    ("mxb_ep_push_counts", _I, [_P, _P, _I, C.POINTER(_P), _I, _I, _P]),
    ("mxb_ep_plan", _I, [_P, _P, _I, _I, _I, C.POINTER(_P), C.POINTER(_P), C.POINTER(_P), _P, _P, _P, _P]),
    ("mxb_quant_act_push", _I, [_P, _P, _I, _P, _I, _P, C.POINTER(MxbActQ), _I, _P, _I, _I, _P]),
    ("mxb_combine", _I, [_P, _P, _I, _P, _P, _I, _I, _P, _P, _P]),
    ("mxb_combine_gated", _I, [_P, _P, _I, _P, _P, _I, _I, _P, _P, _P, _P]),
    ("mxb_quant_pack_weight", _I, [_P, _P, _I, _I, _I, _I, _I, _I, _P, _P, _P]),
    ("mxb_repack_from_ref", _I, [_P, _P, _I, _I, _I, _I, _I, _P, _P]),
    ("mxb_interleave_gate_up", _I, [_P, _P, _I, _I64, _P, _P]),
    ("mxb_packed_bytes", _I64, [_I, _I, _I]),
    ("mxb_scale_elems", _I64, [_I, _I, _I, _I, _I, _I]),
    ("mxb_select_tiles", _I, [_P, C.POINTER(MxbProblem), _I, C.POINTER(MxbTileChoice)]),
]

_lib = None


def load_library(path: str = _LIB_PATH):
    """dlopen the C-ABI library and bind every export; raises if it is missing (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(path):
        raise MxbError(f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                       "(make -C mxmoe_b200/csrc). mxmoe_b200 has no CPU or PyTorch fallback.")
    lib = C.CDLL(path)
    for name, res, args in EXPORTS:
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    for which, st in enumerate((MxbProblem, MxbActQ, MxbRowDest, MxbTileChoice)):
        if lib.mxb_struct_size(which) != C.sizeof(st):
            raise MxbError(f"{path}: sizeof({st.__name__}) is {lib.mxb_struct_size(which)} in the library, {C.sizeof(st)} in the ctypes "
                           "mirror — rebuild the library (make -C mxmoe_b200/csrc)")
    _lib = lib
    return lib


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def permute_static(topk_ids: torch.Tensor, E: int, offs: torch.Tensor, src: torch.Tensor, slot: torch.Tensor):
    """mxb_permute_static: the permute with its scratch in a file-scope device array (no handle, no allocation; at most
    MXB_PERMUTE_STATIC_BYTES of block histograms, see include/mxmoe_b200.h).  Every argument is the caller's: topk_ids
    int32 [T, k] contiguous, offs int32 [E+1], src int32 [>= T*k], slot int32 [T, k].  Calls that overlap on the GPU race."""
    lib = load_library()
    if topk_ids.dtype != torch.int32 or not topk_ids.is_contiguous() or not topk_ids.is_cuda:
        raise MxbError("permute_static takes a contiguous int32 CUDA tensor (it does not convert: converting allocates)")
    T, k = topk_ids.shape
    rc = lib.mxb_permute_static(topk_ids.device.index, _ptr(topk_ids), T, k, E, _ptr(offs), _ptr(src), _ptr(slot), _stream())
    if rc != 0:
        raise MxbError(f"mxb_permute_static failed ({rc}): {lib.mxb_last_error().decode()}")
    return offs, src, slot


@dataclass
class Problem:
    """One GroupGEMM problem in native operand layouts (see include/mxmoe_b200.h)."""
    A: torch.Tensor
    B: torch.Tensor
    C: torch.Tensor
    M: int
    N: int
    K: int
    w_bits: int = 16
    a_bits: int = 16
    gsize: int = -1
    sym: bool = True
    scale_a: Optional[torch.Tensor] = None
    scale_b: Optional[torch.Tensor] = None
    m_range: Optional[torch.Tensor] = None  # device int32 view of {row_begin,row_end}
    lda: int = 0
    ldc: int = 0
    tile_n: int = 0
    flags: int = 0  # EPI_SILU_MUL: B rows interleaved (gate_i, up_i), C = silu(gate)*up [M, N/2]
    c_row: Optional[torch.Tensor] = None  # device int32: first row of the problem in C when it is not m_range[0]
    a_rows: Optional[torch.Tensor] = None  # device int32 per sorted row: the row of A to read (streaming path)
    lds_a: int = 0  # fp16 elements per row of scale_a; 0 = dense (1 per-token, K/128 group-wise)

    def to_c(self) -> MxbProblem:
        return MxbProblem(_ptr(self.A), _ptr(self.B), _ptr(self.scale_a), _ptr(self.scale_b), _ptr(self.C),
                          _ptr(self.m_range), self.lda, self.ldc, self.M, self.N, self.K, self.a_bits, self.w_bits,
                          self.gsize, 1 if self.sym else 0, self.tile_n, self.flags, _ptr(self.c_row), _ptr(self.a_rows), self.lds_a)


class Handle:
    """mxb_handle_t owner.  One per (device, stream-of-use)."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        if not torch.cuda.is_available():
            raise MxbError("mxmoe_b200 needs a CUDA device (sm_100); there is no CPU path")
        h = C.c_void_p()
        rc = self.lib.mxb_create(device, C.byref(h))
        if rc != 0:
            raise MxbError(f"mxb_create failed ({rc}): {self.lib.mxb_last_error().decode()}")
        self.h = h
        self.device = device
        self._keep = {}

    def close(self):
        if getattr(self, "h", None):
            self.lib.mxb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int, what: str):
        if rc != 0:
            raise MxbError(f"{what} failed ({rc}): {self.lib.mxb_last_error().decode()}")

    @property
    def sm_count(self) -> int:
        return self.lib.mxb_sm_count(self.h)

    # ---------------------------------------------------------------- expert-parallel exchange over peer memory
    @staticmethod
    def _ptr_table(tensors):
        if tensors is None:
            return None
        return (C.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])

    def ep_push_counts(self, expert_offsets: torch.Tensor, E: int, peer_counts, rank: int):
        """my per-expert counts -> row `rank` of every rank's [W, E] count matrix (peer_counts: one tensor per rank)"""
        tab = self._ptr_table(peer_counts)
        self._check(self.lib.mxb_ep_push_counts(self.h, expert_offsets.data_ptr(), E, tab, len(peer_counts), rank, _stream()),
                    "mxb_ep_push_counts")

    def ep_plan(self, counts_all: torch.Tensor, W: int, rank: int, E: int, peer_q8, peer_f16, peer_scales, seg_offs: torch.Tensor,
                back_row: torch.Tensor, dest: torch.Tensor):
        """dest: uint8 device tensor of E * sizeof(MxbRowDest) bytes, filled on the device"""
        self._check(self.lib.mxb_ep_plan(self.h, counts_all.data_ptr(), W, rank, E, self._ptr_table(peer_q8), self._ptr_table(peer_f16),
                                         self._ptr_table(peer_scales), seg_offs.data_ptr(), back_row.data_ptr(), dest.data_ptr(),
                                         _stream()), "mxb_ep_plan")

    def quant_act_push(self, x: torch.Tensor, src_row: Optional[torch.Tensor], n_rows: int, expert_offsets: torch.Tensor, actq,
                       dest: torch.Tensor, scale_cols: int, max_ctas: int = 0):
        arr = (MxbActQ * len(actq))(*[MxbActQ(a, g) for a, g in actq])
        self._check(self.lib.mxb_quant_act_push(self.h, x.data_ptr(), x.shape[1], _ptr(src_row), n_rows, expert_offsets.data_ptr(), arr,
                                                len(actq), dest.data_ptr(), scale_cols, max_ctas, _stream()), "mxb_quant_act_push")

    def set_sm_limit(self, sms: int):
        """CTAs of the persistent kernels (default one per SM); <= 0 restores the default."""
        self._check(self.lib.mxb_set_sm_limit(self.h, int(sms)), "mxb_set_sm_limit")

    def set_gemm_path(self, path: int):
        """0 = auto (CTA-pair kernel when the whole list is eligible), 1 = single-CTA kernel, 2 = CTA-pair kernel."""
        self._check(self.lib.mxb_set_gemm_path(self.h, int(path)), "mxb_set_gemm_path")

    # ---------------------------------------------------------------- grouped GEMM
    def pack_problems(self, problems: Sequence[Problem]):
        """Build the ctypes array once; reuse it across calls to hit the plan cache without re-marshalling."""
        arr = (MxbProblem * len(problems))(*[p.to_c() for p in problems])
        return arr

    def group_gemm(self, problems, acc_out: Optional[Sequence[Optional[torch.Tensor]]] = None):
        arr = problems if isinstance(problems, C.Array) else self.pack_problems(problems)
        n = len(arr)
        if acc_out is None:
            rc = self.lib.mxb_group_gemm(self.h, arr, n, _stream())
        else:
            ptrs = (C.c_void_p * n)(*[_ptr(t) for t in acc_out])
            rc = self.lib.mxb_group_gemm_debug_acc(self.h, arr, n, ptrs, _stream())
        self._check(rc, "mxb_group_gemm")

    def interleave_gate_up(self, t: torch.Tensor) -> torch.Tensor:
        """[gate rows; up rows] -> rows (gate_0, up_0, gate_1, up_1, ...) of a packed weight matrix or a scale vector."""
        t = t.contiguous()
        rows = t.shape[0]
        if t.dim() == 2 and rows == 1:  # per-channel scales in the group-major [1, N(,2)] layout of permute_scale
            rows = t.shape[1]
        assert rows % 2 == 0
        row_bytes = t.numel() * t.element_size() // rows
        out = torch.empty_like(t)
        self._check(self.lib.mxb_interleave_gate_up(self.h, _ptr(t), rows // 2, row_bytes, _ptr(out), _stream()),
                    "mxb_interleave_gate_up")
        return out

    def select_tiles(self, problems: Sequence[Problem]) -> List[MxbTileChoice]:
        arr = self.pack_problems(problems)
        out = (MxbTileChoice * len(arr))()
        self._check(self.lib.mxb_select_tiles(self.h, arr, len(arr), out), "mxb_select_tiles")
        return list(out)

    # ---------------------------------------------------------------- MoE pieces
    def permute(self, topk_ids: torch.Tensor, E: int, offs=None, src=None, slot=None):
        """offs / src / slot: optional preallocated outputs (int32 [E+1], [>= T*k], [>= T, k])"""
        T, k = topk_ids.shape
        dev = topk_ids.device
        ids = topk_ids.to(torch.int32).contiguous()
        offs = torch.empty(E + 1, dtype=torch.int32, device=dev) if offs is None else offs
        src = torch.empty(T * k, dtype=torch.int32, device=dev) if src is None else src
        slot = torch.empty(T, k, dtype=torch.int32, device=dev) if slot is None else slot
        self._check(self.lib.mxb_permute(self.h, _ptr(ids), T, k, E, _ptr(offs), _ptr(src), _ptr(slot), _stream()),
                    "mxb_permute")
        return offs, src, slot

    @staticmethod
    def _actq(actq: Sequence[Tuple[int, int]]):
        return (MxbActQ * len(actq))(*[MxbActQ(a, g) for a, g in actq])

    def quant_act(self, x, src_row, n_rows, expert_offsets, actq, hadamard=False, out_q8=None, out_f16=None,
                  scales=None):
        H = x.shape[-1]
        dev = x.device
        any_q = any(a < 16 for a, _ in actq)
        any_f = any(a >= 16 for a, _ in actq)
        cols = H // 128 if any(a < 16 and g == 128 for a, g in actq) else 1
        if any_q and out_q8 is None:
            out_q8 = torch.empty(n_rows, H, dtype=torch.int8, device=dev)
        if any_q and scales is None:
            scales = torch.empty(n_rows, cols, dtype=torch.float16, device=dev)
        if any_f and out_f16 is None:
            out_f16 = torch.empty(n_rows, H, dtype=torch.float16, device=dev)
        self._check(self.lib.mxb_quant_act(self.h, _ptr(x), H, _ptr(src_row), n_rows, _ptr(expert_offsets),
                                           self._actq(actq), len(actq), int(hadamard), _ptr(out_q8), _ptr(out_f16),
                                           _ptr(scales), cols, _stream()), "mxb_quant_act")
        return out_q8, out_f16, scales

    def silu_mul_quant(self, gate_up, n_rows, expert_offsets, actq, hadamard=False, out_q8=None, out_f16=None,
                       scales=None):
        I = gate_up.shape[-1] // 2
        dev = gate_up.device
        any_q = any(a < 16 for a, _ in actq)
        any_f = any(a >= 16 for a, _ in actq)
        cols = I // 128 if any(a < 16 and g == 128 for a, g in actq) else 1
        if any_q and out_q8 is None:
            out_q8 = torch.empty(n_rows, I, dtype=torch.int8, device=dev)
        if any_q and scales is None:
            scales = torch.empty(n_rows, cols, dtype=torch.float16, device=dev)
        if any_f and out_f16 is None:
            out_f16 = torch.empty(n_rows, I, dtype=torch.float16, device=dev)
        self._check(self.lib.mxb_silu_mul_quant(self.h, _ptr(gate_up), I, n_rows, _ptr(expert_offsets),
                                                self._actq(actq), len(actq), int(hadamard), _ptr(out_q8),
                                                _ptr(out_f16), _ptr(scales), cols, _stream()), "mxb_silu_mul_quant")
        return out_q8, out_f16, scales

    def combine(self, down_out, slot_of, topk_w, shared_out=None, out=None, shared_scale=None):
        """shared_scale: optional fp32 [T] multiplying the shared expert's rows (Qwen's sigmoid shared-expert gate)"""
        T, k = slot_of.shape
        H = down_out.shape[-1]
        if out is None:
            out = torch.empty(T, H, dtype=torch.float16, device=down_out.device)
        if shared_scale is not None:
            self._check(self.lib.mxb_combine_gated(self.h, _ptr(down_out), H, _ptr(slot_of), _ptr(topk_w), T, k, _ptr(shared_out),
                                                   _ptr(shared_scale), _ptr(out), _stream()), "mxb_combine_gated")
            return out
        self._check(self.lib.mxb_combine(self.h, _ptr(down_out), H, _ptr(slot_of), _ptr(topk_w), T, k, _ptr(shared_out),
                                         _ptr(out), _stream()), "mxb_combine")
        return out

    # ---------------------------------------------------------------- offline weights
    def quant_pack_weight(self, W: torch.Tensor, w_bits: int, a_bits: int, gsize: int, sym: bool):
        """fp16 [N,K] -> (packed uint8 native layout, scales fp16 group-major).  quantize.cuh:218-475 numerics."""
        N, K = W.shape
        W = W.contiguous()
        packed = torch.empty(self.lib.mxb_packed_bytes(N, K, w_bits), dtype=torch.uint8, device=W.device)
        ne = self.lib.mxb_scale_elems(N, K, w_bits, a_bits, gsize, int(sym))
        scales = torch.empty(ne, dtype=torch.float16, device=W.device)
        self._check(self.lib.mxb_quant_pack_weight(self.h, _ptr(W), N, K, w_bits, a_bits, gsize, int(sym), _ptr(packed),
                                                   _ptr(scales), _stream()), "mxb_quant_pack_weight")
        return packed.view(N, K * abs(w_bits) // 8), scales

    def repack_from_ref(self, ref_packed: torch.Tensor, N: int, K: int, w_bits: int, a_bits: int, sym: bool):
        packed = torch.empty(N, K * w_bits // 8, dtype=torch.uint8, device=ref_packed.device)
        self._check(self.lib.mxb_repack_from_ref(self.h, _ptr(ref_packed), N, K, w_bits, a_bits, int(sym), _ptr(packed),
                                                 _stream()), "mxb_repack_from_ref")
        return packed


_default: Optional[Handle] = None


def default_handle() -> Handle:
    global _default
    if _default is None:
        _default = Handle(torch.cuda.current_device())
    return _default
