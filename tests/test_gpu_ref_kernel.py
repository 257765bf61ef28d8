"""The REFERENCE's own GroupGEMM device code (oracle/_ref/libref_gg.so: its cta_gemm / mm_tile templates compiled in
place for sm_100a by oracle/Makefile) against this library and against the oracle, on identical quantized operands.

What it pins, by EXECUTION of the reference's code rather than by restatement: the integer accumulation, the epilogue
scale order `float(acc) * float(half(sa * sb))` with one rounding to fp16 (mm_tile.cuh:480-496, :636-648) and the
packed operand layout of pack_wxax (quantize.cuh:424-475).  Integer kinds must agree bit for bit three ways
(reference kernel == oracle == mxb_group_gemm); fp16 within 1e-3 of the output range (different fp32 summation order).

A DEFECT of the reference found this way: GlobalTileC::load_scale (mm_tile.cuh:451-464) fetches the per-channel scales of
an 8-column MMA block j at `j*8 + lane%4` and `+1`, while the accumulator fragment of that lane holds columns
`j*8 + (lane%4)*2` and `+1` (its own write_frag_to_gmem, mm_tile.cuh:631).  Lanes with lane%4 != 0 therefore scale their
columns with a NEIGHBOURING channel's scale.  The reference never notices (its quantized check path feeds raw fp16
bytes, test.cu:518-519).  This library and the oracle apply scale_b[n] to column n, as the quantizer defines it
(quant.py:40-106).  Both facts are tested: with scales that are constant inside every aligned 8-channel block the defect
is invisible and the three implementations agree bit for bit; with general scales the reference kernel equals the
oracle evaluated with the mis-indexed scale vector, bit for bit.
Skipped when the reference kernel library was not built (no /root/reference at build time)."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
LIB = os.path.join(ROOT, "oracle", "_ref", "libref_gg.so")


class Dim3(C.Structure):
    _fields_ = [("x", C.c_uint32), ("y", C.c_uint32), ("z", C.c_uint32)]


class QParams(C.Structure):  # quantize.cuh:14-25: int2 qbits (x = a_bits, y = w_bits); int gsize; bool sym
    _fields_ = [("a_bits", C.c_int32), ("w_bits", C.c_int32), ("gsize", C.c_int32), ("sym", C.c_uint8), ("pad", C.c_uint8 * 3)]


def load_ref():
    if not os.path.exists(LIB):
        pytest.skip("oracle/_ref/libref_gg.so not built (needs the reference tree at build time)")
    lib = C.CDLL(LIB)
    lib.ref_gg_launch.restype = C.c_int
    lib.ref_gg_launch.argtypes = [C.c_void_p] * 9 + [C.c_int, C.c_void_p, C.c_int, C.c_void_p]
    assert lib.ref_gg_qparams_size() == C.sizeof(QParams) == 16
    return lib


def ptr_table(tensors):
    return torch.tensor([t.data_ptr() if t is not None else 0 for t in tensors], dtype=torch.int64, device="cuda")


def misindexed(sw):
    """scale_b as the reference's load_scale applies it: column 8j + 2c + d gets the scale of channel 8j + c + d."""
    n = np.arange(sw.shape[0])
    j, c, d = n // 8, (n % 8) // 2, n % 2
    return sw[8 * j + c + d]


def build(rng, shapes, block_scales=True):
    """shapes: list of (M, N, K, bits).  Returns the operands in both layouts + oracle outputs.
    block_scales: every aligned block of 8 weight rows shares its absmax (hence its scale)."""
    items = []
    for (M, N, K, bits) in shapes:
        X = ((rng.random((M, K)) * 2 - 1)).astype(np.float16)
        W = ((rng.random((N, K)) * 2 - 1) * 0.25).astype(np.float16)
        if block_scales:
            W[:, 0] = (0.3 + 0.01 * (np.arange(N) // 8)).astype(np.float16)
        if bits == 16:
            items.append(dict(M=M, N=N, K=K, bits=16, A_ref=X, B_ref=W, A_nat=X, B_nat=W, sa=None, sb=None, ref=O.gemm_fp16(X, W)))
            continue
        qa, sa = O.quant_act(X, bits)
        qw, sw, _ = O.quant_weight(W, bits, -1, True)
        items.append(dict(M=M, N=N, K=K, bits=bits,
                          A_ref=O.pack_wxax(qa, bits), B_ref=O.pack_wxax(qw, bits),          # the reference's packed halves
                          A_nat=qa.astype(np.int8), B_nat=O.pack_wxax_native(qw, bits),       # this library's native layouts
                          sa=np.ascontiguousarray(sa.reshape(M)), sb=np.ascontiguousarray(sw.reshape(N)),
                          ref=O.gemm_wxax(qa, sa, qw, sw, -1)[1],
                          ref_defect=O.gemm_wxax(qa, sa, qw, misindexed(sw.reshape(N)).reshape(N, 1), -1)[1]))
    return items


def run_reference(lib, items):
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda() if a is not None else None
    A = [dev(it["A_ref"].view(np.int16) if it["A_ref"].dtype == np.uint16 else it["A_ref"]) for it in items]
    B = [dev(it["B_ref"].view(np.int16) if it["B_ref"].dtype == np.uint16 else it["B_ref"]) for it in items]
    sa, sb = [dev(it["sa"]) for it in items], [dev(it["sb"]) for it in items]
    Cs = [torch.zeros(it["M"], it["N"], dtype=torch.float16, device="cuda") for it in items]
    P = len(items)
    sizes = (Dim3 * P)(*[Dim3(it["M"], it["N"], it["K"]) for it in items])
    qps = (QParams * P)(*[QParams(it["bits"], it["bits"], -1, 1) for it in items])
    d_sizes = torch.frombuffer(bytearray(bytes(sizes)), dtype=torch.uint8).cuda()
    d_q = torch.frombuffer(bytearray(bytes(qps)), dtype=torch.uint8).cuda()
    d_prefix = torch.zeros(P, dtype=torch.int32, device="cuda")
    tabs = [ptr_table(t) for t in (A, B, sa, sb, Cs)]
    torch.cuda.synchronize()
    rc = lib.ref_gg_launch(*[t.data_ptr() for t in tabs], d_sizes.data_ptr(), C.addressof(sizes), d_q.data_ptr(), C.addressof(qps), P,
                           d_prefix.data_ptr(), 0, None)
    assert rc == 0, f"ref_gg_launch failed: {rc}"
    torch.cuda.synchronize()
    return [c.cpu().numpy() for c in Cs]


def run_ours(items):
    from mxmoe_b200 import ops
    h = ops.Handle(0)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda() if a is not None else None
    probs, outs, keep = [], [], []
    for it in items:
        A, B, sa, sb = dev(it["A_nat"]), dev(it["B_nat"]), dev(it["sa"]), dev(it["sb"])
        Cb = torch.zeros(it["M"], it["N"], dtype=torch.float16, device="cuda")
        keep += [A, B, sa, sb]
        outs.append(Cb)
        if it["bits"] == 16:
            probs.append(ops.Problem(A, B, Cb, it["M"], it["N"], it["K"]))
        else:
            probs.append(ops.Problem(A, B, Cb, it["M"], it["N"], it["K"], it["bits"], it["bits"], -1, True, scale_a=sa, scale_b=sb))
    h.group_gemm(probs)
    torch.cuda.synchronize()
    return [c.cpu().numpy() for c in outs]


def test_reference_kernel_oracle_and_library_agree():
    """One mixed launch of each side: fp16 + w8a8 + w4a4 problems with ragged M (tile tails) and several 128 x 128 tiles."""
    lib = load_ref()
    rng = np.random.default_rng(7)
    items = build(rng, [(200, 256, 256, 16), (300, 384, 512, 8), (130, 256, 512, 4), (128, 128, 1024, 8), (546, 256, 2048, 4)])
    got_ref = run_reference(lib, items)
    got_ours = run_ours(items)
    for it, r, o in zip(items, got_ref, got_ours):
        exp = it["ref"]
        if it["bits"] == 16:
            scale = np.abs(exp.astype(np.float64)).max()
            assert np.abs(r.astype(np.float64) - exp.astype(np.float64)).max() <= 1e-3 * scale
            assert np.abs(o.astype(np.float64) - r.astype(np.float64)).max() <= 1e-3 * scale
        else:
            assert np.array_equal(r.view(np.uint16), exp.view(np.uint16)), f"reference kernel != oracle (w{it['bits']}a{it['bits']})"
            assert np.array_equal(o.view(np.uint16), r.view(np.uint16)), f"mxb_group_gemm != reference kernel (w{it['bits']}a{it['bits']})"


def test_reference_kernel_scale_indexing_defect_is_understood():
    """General per-channel scales: the reference kernel's output is the oracle's with scale_b mis-indexed exactly as
    load_scale does (bit for bit) — and differs from the correctly scaled result, which this library produces."""
    lib = load_ref()
    rng = np.random.default_rng(8)
    items = build(rng, [(300, 384, 512, 8), (130, 256, 512, 4)], block_scales=False)
    got_ref = run_reference(lib, items)
    got_ours = run_ours(items)
    for it, r, o in zip(items, got_ref, got_ours):
        assert np.array_equal(r.view(np.uint16), it["ref_defect"].view(np.uint16))
        assert np.array_equal(o.view(np.uint16), it["ref"].view(np.uint16))
        assert not np.array_equal(r.view(np.uint16), it["ref"].view(np.uint16))
        # columns whose lane has lane % 4 == 0 (columns 0, 1 of every 8) are unaffected
        cols = np.arange(it["N"]) % 8 < 2
        assert np.array_equal(r[:, cols].view(np.uint16), it["ref"][:, cols].view(np.uint16))
