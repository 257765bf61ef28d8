"""CPU oracle for the mixed-precision MoE GroupGEMM path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module; the product (``mxmoe_b200``) never does.

Every function restates a piece of the reference (paths relative to /root/reference):

  quant_minmax / fake_quant     mxmoe/quant/quant.py:40-110
  quant_weight                  mxmoe/kernels/src/include/quantize.cuh:218-279 (zero-scale guard :256)
  pack_wxax / pack_weightonly   mxmoe/kernels/src/include/quantize.cuh:318-475
  permute_scale                 mxmoe/kernels/src/include/quantize.cuh:299-315
  scale order of the epilogue   mxmoe/kernels/src/include/mm_tile.cuh:472-496, 610-662
  group-wise WxAx accumulate    mxmoe/kernels/src/include/cta_gemm.cuh:580-602
  W-only dequant                mxmoe/kernels/src/include/quantize.cuh:134-214 (hfma2 :189-190)
  Hadamard                      mxmoe/quant/hadamard_utils.py:5-105, rotation.py:403-421
  permute                       mxmoe/kernels/src/ref_bind.cu:47-64
  routing weights / combine     mxmoe/quant/moe_tracer.py:79-86

Parity pin status:
  * quant_minmax, Hadamard, workload generation: PINNED against outputs of the reference's
    own Python run in the build container (tests/golden/*.npz, made by tests/golden/gen_golden.py).
  * packers: PINNED against oracle/_ref/ref_pack, compiled from the reference header in place.
  * quantized GroupGEMM end-to-end: the reference has no working check for it
    (test.cu:518-519 feeds raw fp16 bytes) -> "parity unpinned" for that composition;
    it is the composition of the pinned pieces plus the epilogue order restated from mm_tile.cuh.

All half-precision arithmetic is emulated with numpy float16 (each op rounds to nearest even),
which is what torch CPU does for fp16 tensors (compute in fp32, round to fp16).
"""
from __future__ import annotations

import os
import numpy as np

f16, f32, f64 = np.float16, np.float32, np.float64

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "mxmoe_b200", "data")


# --------------------------------------------------------------------------- quantization
def _h_div(a, b):
    """fp16 division as torch-CPU / __fdiv_rn + cvt.rn.f16 do it: fp32 divide, round to fp16."""
    return (a.astype(f32) / b.astype(f32)).astype(f16)


def quant_minmax(t: np.ndarray, bits: int, gsize: int, sym: bool, guard_zero: bool = False):
    """quant.py:40-84 for an fp16 tensor (clip=(1,1)).  Returns (q[int32], scale[f16], zp[f16]).

    q has t's shape; scale/zp have shape [rows*groups, 1] like the reference.
    guard_zero=True applies quantize.cuh:256 (scale==0 -> 1), which the Python path lacks.
    """
    assert t.dtype == f16
    g = t.shape[-1] if gsize == -1 else gsize
    r = t.reshape(-1, g)
    upper = (1 << (bits - 1)) - 1 if sym else (1 << bits) - 1
    lower = -upper if sym else 0
    if not sym:
        gmin = r.min(axis=1, keepdims=True)
        gmax = r.max(axis=1, keepdims=True)
        scale = _h_div((gmax - gmin).astype(f16), np.full_like(gmin, upper))
        zp = gmin
    else:
        gmax = np.abs(r).max(axis=1, keepdims=True)
        scale = _h_div(gmax, np.full_like(gmax, upper))
        zp = np.zeros_like(gmax)
    if guard_zero:
        scale = np.where(scale == 0, f16(1), scale).astype(f16)
    with np.errstate(divide="ignore", invalid="ignore"):
        q = _h_div((r - zp).astype(f16), scale)
    q = np.rint(np.clip(q, f16(lower), f16(upper)))
    return q.reshape(t.shape).astype(np.int32), scale, zp


def dequant(q: np.ndarray, scale, zp, gsize: int):
    """quant.py:99-102 in fp16."""
    g = q.shape[-1] if gsize == -1 else gsize
    r = q.reshape(-1, g).astype(f16)
    return ((r * scale).astype(f16) + zp).astype(f16).reshape(q.shape)


def fake_quant(t, bits, gsize, sym):
    if bits == 16:
        return t
    if bits == BITS_E4M3:  # per-row E4M3 grid (defined below): value * scale
        v, s = quant_e4m3(np.ascontiguousarray(t, dtype=f16))
        return (v * s.astype(f64)).astype(f32)
    q, s, z = quant_minmax(t, bits, gsize, sym)
    return dequant(q, s, z, gsize)


def quant_weight(W: np.ndarray, bits: int, gsize: int, sym: bool):
    """quantize.cuh:218-279: per (row, group) min/max in fp16, scale==0 -> 1.

    Returns q int32 [N,K], scale f16 [N, K/g], zp f16 [N, K/g] (zeros when sym).
    """
    N, K = W.shape
    g = K if gsize == -1 else gsize
    q, s, z = quant_minmax(W, bits, gsize, sym, guard_zero=True)
    return q, s.reshape(N, K // g), z.reshape(N, K // g)


def quant_act(X: np.ndarray, bits: int, gsize: int = -1):
    """Per-token symmetric activation quantization (quant.py:126-130 hook -> quant_minmax, sym).

    Returns q int32 [M,K], scale f16 [M, K/g].  All-zero rows get scale 1 (kernel guard).
    """
    M, K = X.shape
    g = K if gsize == -1 else gsize
    q, s, _ = quant_minmax(X, bits, gsize, True, guard_zero=True)
    return q, s.reshape(M, K // g)


# --------------------------------------------------------------------------- 8-bit float (E4M3) grids
# The reference only NAMES this kind (`w8a8_g-1_sym_E4M3`, mxmoe/kernels/tile_config.py:45,105,192: MMA_E4M3_K32,
# USE_FP=True) — it has no quantizer, packer or test for it, so this restatement is "parity unpinned": the definition
# below is the int path's (quant.py:40-84) with the integer grid replaced by the E4M3 one:
#   scale = fp16(absmax / 448) (0 -> 1), code = e4m3_rne_satfinite(x / scale), value = code * scale.
BITS_E4M3 = -8
E4M3_MAX = 448.0


def _e4m3_grid():
    v = [0.0]
    for e in range(16):
        for m in range(8):
            if e == 15 and m == 7:
                continue  # NaN
            v.append((m / 8.0) * 2.0 ** -6 if e == 0 else (1 + m / 8.0) * 2.0 ** (e - 7))
    g = np.unique(np.array(v, f64))
    return g


_E4M3 = _e4m3_grid()


def e4m3_round(x: np.ndarray) -> np.ndarray:
    """Round to the nearest E4M3 value (ties to the even mantissa), saturating at +-448.  Returns float64 values."""
    a = np.minimum(np.abs(x.astype(f64)), E4M3_MAX)
    hi = np.searchsorted(_E4M3, a, side="left").clip(0, len(_E4M3) - 1)
    lo = (hi - 1).clip(0)
    dlo, dhi = a - _E4M3[lo], _E4M3[hi] - a
    # grid index parity == mantissa parity (the grid lists every code in order, 8 mantissas per binade)
    pick_hi = (dhi < dlo) | ((dhi == dlo) & (hi % 2 == 0))
    r = np.where(pick_hi, _E4M3[hi], _E4M3[lo])
    return np.copysign(r, x)


def e4m3_encode(v: np.ndarray) -> np.ndarray:
    """E4M3 values (as produced by e4m3_round) -> byte codes (sign | exponent | mantissa)."""
    a = np.abs(v.astype(f64))
    idx = np.searchsorted(_E4M3, a)  # grid index == code of the magnitude (0, subnormals 1..7, then 8 per binade)
    return (idx.astype(np.uint8) | (np.signbit(v).astype(np.uint8) << 7)).astype(np.uint8)  # -0 keeps its sign, like cvt.rn.satfinite.e4m3x2


def quant_e4m3(t: np.ndarray):
    """Per-row E4M3 quantization of an fp16 [R, K] tensor: returns (values float64 [R,K], scale f16 [R,1])."""
    assert t.dtype == f16
    amax = np.abs(t).max(axis=1, keepdims=True)
    scale = _h_div(amax, np.full_like(amax, E4M3_MAX))
    scale = np.where(scale == 0, f16(1), scale).astype(f16)
    return e4m3_round(t.astype(f64) / scale.astype(f64)), scale


def gemm_e4m3(vA, sA, vW, sW):
    """kind::f8f6f4 problem: fp32 accumulation of exact E4M3 products (here float64), then the int path's scale order
    out = fp16( acc * fp32(fp16(sA * sW)) ) (mm_tile.cuh:472-496)."""
    acc = (vA @ vW.T).astype(f32)
    prod = (sA.astype(f16) * sW.astype(f16).reshape(1, -1)).astype(f16)
    return (acc.astype(f64) * prod.astype(f64)).astype(f32).astype(f16)


# --------------------------------------------------------------------------- packed layouts
def pack_wxax(q: np.ndarray, bits: int) -> np.ndarray:
    """quantize.cuh:424-475: two's-complement values, PACK=16/bits per uint16 along K,
    first element in the most-significant bits.  Returns uint16 [rows, K/PACK]."""
    assert bits in (4, 8)
    pack = 16 // bits
    R, K = q.shape
    assert K % pack == 0
    v = (q.astype(np.int64) & ((1 << bits) - 1)).astype(np.uint16).reshape(R, K // pack, pack)
    out = np.zeros((R, K // pack), np.uint16)
    for x in range(pack):
        out = (out << np.uint16(bits)) | v[:, :, x]
    return out


def unpack_wxax(p: np.ndarray, bits: int) -> np.ndarray:
    pack = 16 // bits
    R, C = p.shape
    out = np.zeros((R, C, pack), np.int32)
    for x in range(pack):
        sh = bits * (pack - 1 - x)
        v = ((p >> np.uint16(sh)) & np.uint16((1 << bits) - 1)).astype(np.int32)
        out[:, :, x] = np.where(v >= (1 << (bits - 1)), v - (1 << bits), v)
    return out.reshape(R, C * pack)


def permute_scale(scale: np.ndarray, zp, sym: bool) -> np.ndarray:
    """quantize.cuh:299-315: [N, K/g] -> group-major [K/g, N] (asym: interleaved (scale, zp))."""
    if sym:
        return np.ascontiguousarray(scale.T)
    return np.ascontiguousarray(np.stack([scale.T, zp.T], axis=-1))


def _compose_perm(proj, desired, n, plen):
    perm = [0] * n
    for i in range(0, n, plen):
        for j in range(plen):
            perm[proj[j] + i] = desired[i + j]
    return perm


def permute_weight_ref(q: np.ndarray, bits: int) -> np.ndarray:
    """quantize.cuh:318-381 (PermuteMode::Row), vectorised: inside every block of
    (PACK*8 rows) x (2 k) the 2*PACK values of column-in-fragment `trav_j` are shuffled by
    `intermediate_perm` so that the LOP3 dequant order matches the ldmatrix fragment order."""
    N, K = q.shape
    pack = 16 // bits
    if bits == 8:
        perm = _compose_perm([1, 0, 3, 2], [0, 2, 4, 6, 1, 3, 5, 7], 8, 4)
    elif bits == 4:
        perm = _compose_perm([3, 7, 2, 6, 1, 5, 0, 4],
                             [0, 4, 8, 12, 1, 5, 9, 13, 2, 6, 10, 14, 3, 7, 11, 15], 16, 8)
    elif bits == 2:
        perm = _compose_perm([7, 15, 6, 14, 5, 13, 4, 12, 3, 11, 2, 10, 1, 9, 0, 8],
                             [0, 8, 16, 24, 1, 9, 17, 25, 2, 10, 18, 26, 3, 11, 19, 27,
                              4, 12, 20, 28, 5, 13, 21, 29, 6, 14, 22, 30, 7, 15, 23, 31], 32, 16)
    else:
        raise ValueError("not implemented")
    assert N % (pack * 8) == 0 and K % 16 == 0
    # buffer order in the reference: for ii in (0, 8): for trav_ii in (0, 1): for frag in range(pack)
    # element = w[(j + frag*8 + trav_j), i + ii + trav_i + trav_ii]
    # view q as [Nb, frag, trav_j, Kb(16), ii(2), trav_i(4), trav_ii(2)]
    v = q.reshape(N // (pack * 8), pack, 8, K // 16, 2, 4, 2)
    # bring buffer axes (ii, trav_ii, frag) last in that nesting order
    b = v.transpose(0, 2, 3, 5, 4, 6, 1)          # [Nb, trav_j, Kb, trav_i, ii, trav_ii, frag]
    shp = b.shape
    flat = b.reshape(*shp[:4], 4 * pack)
    flat = flat[..., perm]                          # res[ptr[x]] = val[perm[x]]
    b2 = flat.reshape(shp)
    out = b2.transpose(0, 6, 1, 2, 4, 3, 5)        # back to [Nb, frag, trav_j, Kb, ii, trav_i, trav_ii]
    return np.ascontiguousarray(out).reshape(N, K)


def pack_weightonly_ref(q: np.ndarray, bits: int, sym: bool) -> np.ndarray:
    """permute_weight + pack_weightonly (quantize.cuh:318-421).  q: int32 [N,K] (sym: signed,
    asym: unsigned).  Returns uint16 [N/PACK, K]: one word = the same k for PACK rows
    n = j + frag*8 + jj, first fragment in the most-significant bits; sym values biased by
    2^(b-1)-1."""
    N, K = q.shape
    pack = 16 // bits
    p = permute_weight_ref(q, bits)
    if sym:
        p = p + ((1 << (bits - 1)) - 1)
    v = (p.astype(np.int64) & 0xFFFF).astype(np.uint16).reshape(N // (pack * 8), pack, 8, K)
    out = np.zeros((N // (pack * 8), 8, K), np.uint16)
    for f in range(pack):
        out = (out << np.uint16(bits)) | v[:, f]
    return out.reshape(N // pack, K)


def unpack_weightonly_ref(p: np.ndarray, bits: int, sym: bool, N: int, K: int) -> np.ndarray:
    """Inverse of pack_weightonly_ref (used to check mxb_repack_from_ref)."""
    pack = 16 // bits
    w = p.reshape(N // (pack * 8), 8, K)
    v = np.zeros((N // (pack * 8), pack, 8, K), np.int32)
    for f in range(pack):
        sh = bits * (pack - 1 - f)
        v[:, f] = ((w >> np.uint16(sh)) & np.uint16((1 << bits) - 1)).astype(np.int32)
    permuted = v.reshape(N, K)
    if sym:
        permuted = permuted - ((1 << (bits - 1)) - 1)
    # invert permute_weight_ref by applying it to an index matrix
    idx = np.arange(N * K, dtype=np.int32).reshape(N, K)
    pidx = permute_weight_ref(idx, bits)
    out = np.zeros(N * K, np.int32)
    out[pidx.reshape(-1)] = permuted.reshape(-1)
    return out.reshape(N, K)


# --------------------------------------------------------------------------- native W-only layout
def pack_weightonly_native(q: np.ndarray, bits: int, sym: bool) -> np.ndarray:
    """Native (B200) W-only layout consumed by the sm_100a kernels: K-major uint8 [N, K*bits/8].
    Values are the same biased-unsigned codes the reference packs (sym: q + 2^(b-1)-1).
    4-bit: inside every 32-bit word the 8 nibbles (lsb first) hold k = 0,2,4,6,1,3,5,7 so one
    LOP3 per shift yields two adjacent k as a half2.  8-bit: plain K order."""
    N, K = q.shape
    u = q + ((1 << (bits - 1)) - 1) if sym else q
    u = u.astype(np.uint32)
    if bits == 8:
        return u.astype(np.uint8)
    assert bits == 4 and K % 8 == 0
    g = u.reshape(N, K // 8, 8)
    order = [0, 2, 4, 6, 1, 3, 5, 7]
    word = np.zeros((N, K // 8), np.uint32)
    for pos, k in enumerate(order):
        word |= g[:, :, k] << np.uint32(4 * pos)
    return word.view(np.uint8).reshape(N, K // 2)


def pack_wxax_native(q: np.ndarray, bits: int) -> np.ndarray:
    """Native (B200) WxAx weight layout: 8-bit = int8 two's complement [N,K] in K order; 4-bit =
    [N,K/2] bytes, inside every 32-k block byte j holds k=j in the low and k=j+16 in the high nibble
    (two's complement), so the in-smem unpack is one shift/mask per 32-bit word."""
    N, K = q.shape
    if bits == 8:
        return q.astype(np.int8).view(np.uint8)
    assert bits == 4 and K % 32 == 0
    u = (q.astype(np.int64) & 0xF).astype(np.uint8).reshape(N, K // 32, 2, 16)
    return (u[:, :, 0, :] | (u[:, :, 1, :] << np.uint8(4))).reshape(N, K // 2)


# --------------------------------------------------------------------------- GEMM problems
def gemm_fp16(A: np.ndarray, W: np.ndarray) -> np.ndarray:
    """cta_gemm.cuh:8-107: fp16 operands, fp32 accumulate, one rounding to fp16."""
    return (A.astype(f64) @ W.astype(f64).T).astype(f32).astype(f16)


def gemm_wxax(qA, sA, qW, sW, gsize: int):
    """WxAx problem (cta_gemm.cuh:424-608 + mm_tile.cuh:472-496,643).

    qA int [M,K], sA f16 [M,K/g]; qW int [N,K], sW f16 [N,K/g].
    Returns (acc int32 [G,M,N], out f16 [M,N]) with
      out = fp16( sum_g fp32(acc_g) * fp32( fp16(sA_g * sW_g) ) )   accumulated in fp32.
    """
    M, K = qA.shape
    N = qW.shape[0]
    g = K if gsize == -1 else gsize
    G = K // g
    acc = np.zeros((G, M, N), np.int32)
    out = np.zeros((M, N), f32)
    for i in range(G):
        a = qA[:, i * g:(i + 1) * g].astype(f64)
        w = qW[:, i * g:(i + 1) * g].astype(f64)
        acc[i] = np.rint(a @ w.T).astype(np.int32)          # exact: |sum| < 2^53
        prod = (sA[:, i:i + 1].astype(f16) * sW[:, i].astype(f16)[None, :]).astype(f16)
        # fused multiply-add in fp32 (nvcc contracts `out += float(acc) * float(prod)`)
        # float(acc) rounds to fp32 first (|acc| can exceed 2^24), then one fp32 FMA
        out = (acc[i].astype(f32).astype(f64) * prod.astype(f64) + out.astype(f64)).astype(f32)
    return acc, out.astype(f16)


def dequant_wonly(qW, scale, zp, gsize: int, sym: bool) -> np.ndarray:
    """quantize.cuh:189-190: w = hfma2(half(q), scale, zp) — one fp16 rounding."""
    N, K = qW.shape
    g = K if gsize == -1 else gsize
    s = np.repeat(scale.astype(f64), g, axis=1)
    z = np.repeat(zp.astype(f64), g, axis=1) if not sym else 0.0
    return (qW.astype(f64) * s + z).astype(f16)


def gemm_wonly(A, qW, scale, zp, gsize: int, sym: bool) -> np.ndarray:
    """W-only problem (cta_gemm.cuh:113-421): dequant to fp16, fp16 MMA, fp32 accumulate."""
    return gemm_fp16(A, dequant_wonly(qW, scale, zp, gsize, sym))


# --------------------------------------------------------------------------- Hadamard
_HADK = None


def hadK(k: int) -> np.ndarray:
    """The K x K +-1 matrices of hadamard_utils.py:160-4379, shipped as sign bits in
    mxmoe_b200/data/hadk.npz (made by tests/golden/gen_golden.py from the reference)."""
    global _HADK
    if _HADK is None:
        _HADK = np.load(os.path.join(_DATA, "hadk.npz"))
    bits = np.unpackbits(_HADK[f"had{k}"])[: k * k].reshape(k, k)
    return bits.astype(f64) * 2 - 1


HADK_ORDER = (176, 172, 156, 140, 108, 60, 52, 36, 28, 40, 20, 12)


def get_hadK(n: int):
    """hadamard_utils.py:5-59 — same test order (28 before 40, 40 before 20)."""
    for k in HADK_ORDER:
        if n % k == 0:
            m = n // k
            assert m & (m - 1) == 0
            return hadK(k), k
    assert n & (n - 1) == 0
    return None, 1


def hadamard(X: np.ndarray) -> np.ndarray:
    """matmul_hadU (hadamard_utils.py:62-83) == matmul_hadU_cuda (:96-105) in exact arithmetic:
    view (.., K, n/K), Sylvester FWHT over the last axis, hadK @ over the K axis, / sqrt(n).
    Computed in float64; returns float64."""
    n = X.shape[-1]
    hk, k = get_hadK(n)
    x = X.astype(f64).reshape(-1, k, n // k)
    m = n // k
    h = 1
    while h < m:
        x = x.reshape(-1, k, m // (2 * h), 2, h)
        a, b = x[:, :, :, 0, :], x[:, :, :, 1, :]
        x = np.stack([a + b, a - b], axis=3)
        h *= 2
    x = x.reshape(-1, k, m)
    if k > 1:
        x = np.einsum("ij,bjm->bim", hk, x)
    return (x / np.sqrt(f64(n))).reshape(X.shape)


def hadamard_f16(X: np.ndarray) -> np.ndarray:
    """What the online hook produces for an fp16 input (rotation.py:410-414): the FWHT
    result is rounded to fp16 before the hadK GEMM, whose output is rounded again."""
    n = X.shape[-1]
    hk, k = get_hadK(n)
    if k == 1:
        return hadamard(X).astype(f16)
    x = X.astype(f64).reshape(-1, k, n // k)
    m = n // k
    y = hadamard(x.reshape(-1, m)) * np.sqrt(f64(m)) / np.sqrt(f64(n))   # FWHT scaled by 1/sqrt(n)
    y = y.astype(f16).astype(f64).reshape(-1, k, m)
    return np.einsum("ij,bjm->bim", hk, y).astype(f32).astype(f16).reshape(X.shape)


# --------------------------------------------------------------------------- layer pieces
def permute(topk_ids: np.ndarray, E: int):
    """ref_bind.cu:47-64: stable sort of the flattened expert ids.
    Returns (expert_offsets[E+1], src_row[T*k] = source token of each sorted slot,
             slot_of[T,k] = sorted position of (t, j))."""
    T, k = topk_ids.shape
    flat = topk_ids.reshape(-1)
    order = np.argsort(flat, kind="stable")
    counts = np.bincount(flat, minlength=E)
    offs = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
    slot_of = np.empty(T * k, np.int32)
    slot_of[order] = np.arange(T * k, dtype=np.int32)
    return offs, (order // k).astype(np.int32), slot_of.reshape(T, k)


def silu_mul(GU: np.ndarray) -> np.ndarray:
    """act_fn(gate) * up as torch does on fp16: silu in fp32 rounded to fp16, product rounded
    to fp16.  GU = [M, 2I] with gate in the first I columns (gate_up fused, gen_workload.py:94)."""
    I = GU.shape[1] // 2
    g = GU[:, :I].astype(f32)
    s = (g / (f32(1) + np.exp(-g))).astype(f16)
    return (s.astype(f32) * GU[:, I:].astype(f32)).astype(f16)


def combine(down_out: np.ndarray, slot_of: np.ndarray, topk_w: np.ndarray, shared=None) -> np.ndarray:
    """out[t] = sum_j w[t,j] * down_out[slot_of[t,j]] (+ shared[t]); fp32 accumulate in j
    order, one rounding to fp16 (moe_tracer.py:79-86 defines w)."""
    T, k = slot_of.shape
    acc = np.zeros((T, down_out.shape[1]), f32)
    for j in range(k):
        acc = (acc.astype(f64) + topk_w[:, j:j + 1].astype(f64) * down_out[slot_of[:, j]].astype(f64)).astype(f32)
    if shared is not None:
        acc = acc + shared.astype(f32)
    return acc.astype(f16)


def routing_weights(logits: np.ndarray, k: int, renorm: bool):
    """moe_tracer.py:79-86: softmax in fp32 -> top-k -> optional renormalisation."""
    x = logits.astype(f32)
    p = np.exp(x - x.max(axis=1, keepdims=True))
    p /= p.sum(axis=1, keepdims=True)
    ids = np.argsort(-p, axis=1, kind="stable")[:, :k]
    w = np.take_along_axis(p, ids, axis=1)
    if renorm:
        w = w / w.sum(axis=1, keepdims=True)
    return ids.astype(np.int32), w.astype(f32)


def run_problem(A: np.ndarray, W: np.ndarray, w_bits, a_bits, gsize, sym):
    """One GroupGEMM problem from fp16 operands, by qconfig: returns fp16 [M,N]."""
    if w_bits == 16 and a_bits == 16:
        return gemm_fp16(A, W)
    if a_bits == 16:
        q, s, z = quant_weight(W, w_bits, gsize, sym)
        return gemm_wonly(A, q, s, z, gsize, sym)
    if a_bits == BITS_E4M3:
        va, sa = quant_e4m3(A)
        vw, sw = quant_e4m3(W)
        return gemm_e4m3(va, sa, vw, sw)
    qa, sa = quant_act(A, a_bits, gsize)
    qw, sw, _ = quant_weight(W, w_bits, gsize, True)
    return gemm_wxax(qa, sa, qw, sw, gsize)[1]


# --------------------------------------------------------------------------- whole layer
def expert_mlp(X, Wgu, Wdn, q_gu, q_dn, hadamard_down: bool):
    """One expert: gate_up GEMM -> silu*mul -> (online Hadamard on the down input, rotation.py:403-421) -> down GEMM.
    q_* = (w_bits, a_bits, gsize, sym).  Activations are quantized per token inside run_problem."""
    gu = run_problem(X, Wgu, *q_gu)
    h = silu_mul(gu)
    if hadamard_down:
        h = hadamard_f16(h)
    return run_problem(h, Wdn, *q_dn)


def moe_layer(x, topk_ids, topk_w, weights, qcfgs, E, hadamard_down=False, shared=None):
    """The hot path end to end.  weights[e] = (Wgu [2I,H], Wdn [H,I]); qcfgs[e] = (q_gu, q_dn);
    shared = (Wgu, Wdn, q_gu, q_dn) or None.  Returns fp16 [T,H]."""
    offs, src, slot = permute(topk_ids, E)
    xs = x[src]
    Hd = x.shape[1]
    down = np.zeros((len(src), Hd), f16)
    for e in range(E):
        r = slice(offs[e], offs[e + 1])
        if offs[e] == offs[e + 1]:
            continue
        down[r] = expert_mlp(xs[r], weights[e][0], weights[e][1], qcfgs[e][0], qcfgs[e][1], hadamard_down)
    sh = None
    if shared is not None:
        sh = expert_mlp(x, shared[0], shared[1], shared[2], shared[3], hadamard_down)
    return combine(down, slot, topk_w, sh)
