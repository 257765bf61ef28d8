"""Parity of the permute / act-quant (+Hadamard) / silu-mul-quant / combine / weight-pack kernels
(through the C-ABI) against the CPU oracle.  Integer and index outputs: bit-exact.  Paths that go
through a transcendental (silu) or an fp32 butterfly (Hadamard) may land on the other side of an
fp16 rounding boundary in rare elements: the test states the tolerance where that applies."""
import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def H():
    from mxmoe_b200 import ops
    return ops.Handle(0)


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def rand16(rng, *shape, scale=1.0):
    return ((rng.random(shape) * 2 - 1) * scale).astype(np.float16)


@pytest.mark.parametrize("T,k,E", [(1, 1, 1), (37, 4, 12), (512, 4, 60), (8192, 4, 60), (300, 6, 64), (5000, 8, 64), (64, 2, 8)])
def test_permute_bit_exact(H, T, k, E):
    rng = np.random.default_rng(T + E)
    ids = np.stack([rng.permutation(E)[:k] for _ in range(T)]).astype(np.int32)
    offs, src, slot = H.permute(dev(ids), E)
    torch.cuda.synchronize()
    r_offs, r_src, r_slot = O.permute(ids, E)
    assert np.array_equal(offs.cpu().numpy(), r_offs)
    assert np.array_equal(src.cpu().numpy(), r_src)
    assert np.array_equal(slot.cpu().numpy(), r_slot)
    # second call re-uses the workspace ticket
    offs2, _, _ = H.permute(dev(ids), E)
    assert np.array_equal(offs2.cpu().numpy(), r_offs)


def test_permute_skewed_and_empty_experts(H):
    rng = np.random.default_rng(3)
    T, k, E = 4096, 2, 16
    ids = rng.choice([0, 3, 15], size=(T, k)).astype(np.int32)  # 13 experts get nothing, collisions inside a token allowed
    offs, src, slot = H.permute(dev(ids), E)
    r = O.permute(ids, E)
    assert np.array_equal(offs.cpu().numpy(), r[0]) and np.array_equal(src.cpu().numpy(), r[1]) and np.array_equal(slot.cpu().numpy(), r[2])


@pytest.mark.parametrize("Hdim", [2048, 1408, 256, 4096])
def test_quant_act_mixed_experts_bit_exact(H, Hdim):
    rng = np.random.default_rng(Hdim)
    T, k, E = 96, 2, 6
    x = rand16(rng, T, Hdim, scale=3.0)
    x[5] = 0  # all-zero token: scale guard (quantize.cuh:256)
    ids = np.stack([rng.permutation(E)[:k] for _ in range(T)]).astype(np.int32)
    offs, src, _ = O.permute(ids, E)
    actq = [(8, -1), (4, -1), (16, -1), (4, 128), (8, 128), (4, -1)]
    q8, f16, sc = H.quant_act(dev(x), dev(src), T * k, dev(offs), actq)
    torch.cuda.synchronize()
    q8, f16, sc = q8.cpu().numpy(), f16.cpu().numpy(), sc.cpu().numpy()
    xs = x[src]
    for e, (a, g) in enumerate(actq):
        rows = slice(offs[e], offs[e + 1])
        if a == 16:
            assert np.array_equal(f16[rows].view(np.uint16), xs[rows].view(np.uint16))
            continue
        rq, rs = O.quant_act(xs[rows], a, g)
        assert np.array_equal(q8[rows].astype(np.int32), rq), f"expert {e} q"
        assert np.array_equal(sc[rows, :rs.shape[1]].view(np.uint16), rs.view(np.uint16)), f"expert {e} scale"


def test_silu_mul_quant(H):
    rng = np.random.default_rng(11)
    rows, I = 200, 1408
    gu = rand16(rng, rows, 2 * I, scale=4.0)
    offs = np.array([0, 50, 50, 120, 200], np.int32)
    actq = [(8, -1), (4, -1), (4, -1), (16, -1)]
    q8, f16, sc = H.silu_mul_quant(dev(gu), rows, dev(offs), actq)
    torch.cuda.synchronize()
    q8, f16, sc = q8.cpu().numpy(), f16.cpu().numpy(), sc.cpu().numpy()
    h = O.silu_mul(gu)
    # ex2.approx+rcp.approx on the device vs numpy exp: identical after the fp16 rounding except at rounding
    # boundaries (expected ~5e-4 of the elements, one fp16 ulp)
    got = f16[120:200].astype(np.float32)
    ref = h[120:200].astype(np.float32)
    assert np.mean(got != ref) < 5e-3 and np.abs(got - ref).max() <= 2 ** -8 * np.abs(ref).max()
    for e, (a, g) in enumerate(actq[:3]):
        r = slice(offs[e], offs[e + 1])
        if offs[e] == offs[e + 1]:
            continue
        rq, rs = O.quant_act(h[r], a, g)
        # tolerance: a 1-ulp difference in h can move a code by one step and, if it is the row maximum, the scale by 1 ulp
        assert np.mean(q8[r].astype(np.int32) != rq) < 1e-2
        assert np.abs(q8[r].astype(np.int32) - rq).max() <= 1
        assert np.abs(sc[r, 0].astype(np.float32) - rs[:, 0].astype(np.float32)).max() <= 2 ** -10 * np.abs(rs.astype(np.float32)).max()


@pytest.mark.parametrize("n", [2048, 1408, 5632, 2560, 256, 96])
def test_hadamard_quant_fp16_passthrough(H, n):
    """hadamard=1 with a_bits=16 returns the rotated row itself: compare with hadamard_f16 (hadamard_utils.py:96-105).
    Tolerance: fp32 butterflies vs float64 oracle -> the fp16 results may differ by one ulp in rare elements."""
    rng = np.random.default_rng(n)
    rows = 40
    x = rand16(rng, rows, n, scale=2.0)
    x[3, 7] = 40.0  # an outlier the rotation spreads
    _, f16, _ = H.quant_act(dev(x), None, rows, None, [(16, -1)], hadamard=True)
    torch.cuda.synchronize()
    got = f16.cpu().numpy().astype(np.float64)
    ref = O.hadamard_f16(x).astype(np.float64)
    scale = np.abs(ref).max()
    assert np.abs(got - ref).max() <= 2e-3 * scale
    assert np.mean(got != ref) < 0.02
    # orthogonality: energy is preserved
    np.testing.assert_allclose((got ** 2).sum(1), (x.astype(np.float64) ** 2).sum(1), rtol=5e-3)


def test_hadamard_silu_quant_close_to_oracle(H):
    rng = np.random.default_rng(2)
    rows, I = 64, 1408
    gu = rand16(rng, rows, 2 * I, scale=3.0)
    q8, _, sc = H.silu_mul_quant(dev(gu), rows, None, [(8, -1)], hadamard=True)
    torch.cuda.synchronize()
    h = O.hadamard_f16(O.silu_mul(gu))
    rq, rs = O.quant_act(h, 8, -1)
    deq = q8.cpu().numpy().astype(np.float32) * sc.cpu().numpy().astype(np.float32)
    rdeq = rq.astype(np.float32) * rs.astype(np.float32)
    # dequantized rows agree to one quantization step
    assert np.abs(deq - rdeq).max() <= 1.01 * rs.astype(np.float32).max()
    assert np.mean(q8.cpu().numpy().astype(np.int32) != rq) < 0.02


@pytest.mark.parametrize("n,silu,bits", [(20480, False, 8), (10240, True, 4), (20480, True, 16), (7168, False, 8)])
def test_hadamard_long_rows_cta_per_row_kernel(H, n, silu, bits, monkeypatch):
    """Rows whose power-of-two part exceeds 128 (Qwen2-57B's fused shared expert: 20480 = 40 x 512; 7168 = 28 x 256)
    take hadamard_wide_kernel (one CTA per row): against the oracle, and bit for bit against the warp-per-row kernel
    it replaces (MXB_HAD_WIDE=0) -- same butterfly network, same fp16 staging, same MMA tiles, same quant8."""
    rng = np.random.default_rng(n + bits)
    rows = 21
    x = rand16(rng, rows, 2 * n if silu else n, scale=2.0)
    x[2, 5] = 30.0
    actq = [(bits, -1)]

    def run():
        if silu:
            return H.silu_mul_quant(dev(x), rows, None, actq, hadamard=True)
        return H.quant_act(dev(x), None, rows, None, actq, hadamard=True)
    q8, f16, sc = run()
    torch.cuda.synchronize()
    monkeypatch.setenv("MXB_HAD_WIDE", "0")
    q8b, f16b, scb = run()
    torch.cuda.synchronize()
    monkeypatch.delenv("MXB_HAD_WIDE")
    h = O.hadamard_f16(O.silu_mul(x) if silu else x)
    if bits == 16:
        assert torch.equal(f16.view(torch.int16), f16b.view(torch.int16))
        got, ref = f16.cpu().numpy().astype(np.float64), h.astype(np.float64)
        assert np.abs(got - ref).max() <= 2e-3 * np.abs(ref).max()
    else:
        assert torch.equal(q8, q8b) and torch.equal(sc.view(torch.int16), scb.view(torch.int16))
        rq, rs = O.quant_act(h, bits, -1)
        deq = q8.cpu().numpy().astype(np.float32) * sc.cpu().numpy().astype(np.float32)
        assert np.abs(deq - rq.astype(np.float32) * rs.astype(np.float32)).max() <= 1.01 * rs.astype(np.float32).max()
        assert np.mean(q8.cpu().numpy().astype(np.int32) != rq) < 0.02


def test_combine_bit_exact(H):
    rng = np.random.default_rng(4)
    T, k, E, Hd = 333, 4, 12, 2048
    ids = np.stack([rng.permutation(E)[:k] for _ in range(T)]).astype(np.int32)
    _, _, slot = O.permute(ids, E)
    down = rand16(rng, T * k, Hd, scale=2.0)
    w = rng.random((T, k)).astype(np.float32)
    shared = rand16(rng, T, Hd)
    out = H.combine(dev(down), dev(slot), dev(w), dev(shared))
    out2 = H.combine(dev(down), dev(slot), dev(w))
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy().view(np.uint16), O.combine(down, slot, w, shared).view(np.uint16))
    assert np.array_equal(out2.cpu().numpy().view(np.uint16), O.combine(down, slot, w).view(np.uint16))


@pytest.mark.parametrize("w_bits,a_bits,gsize,sym", [(8, 8, -1, True), (4, 4, -1, True), (4, 4, 128, True), (4, 16, 128, False),
                                                     (4, 16, -1, True), (8, 16, 128, True), (8, 16, -1, False), (4, 16, 128, True)])
def test_quant_pack_weight_bit_exact(H, w_bits, a_bits, gsize, sym):
    rng = np.random.default_rng(w_bits + a_bits)
    N, K = 96, 512
    W = rand16(rng, N, K)
    W[7] = 0
    packed, scales = H.quant_pack_weight(dev(W), w_bits, a_bits, gsize, sym)
    torch.cuda.synchronize()
    q, s, z = O.quant_weight(W, w_bits, gsize, sym)
    ref_p = O.pack_wxax_native(q, w_bits) if a_bits < 16 else O.pack_weightonly_native(q, w_bits, sym)
    ref_s = O.permute_scale(s, None if sym else z, sym)
    assert np.array_equal(packed.cpu().numpy(), ref_p.reshape(N, -1))
    assert np.array_equal(scales.cpu().numpy().view(np.uint16), ref_s.reshape(-1).view(np.uint16))


@pytest.mark.parametrize("w_bits,a_bits,sym", [(8, 8, True), (4, 4, True), (4, 16, False), (4, 16, True), (8, 16, True)])
def test_repack_from_reference_layout(H, w_bits, a_bits, sym):
    """Weights packed by the reference's own packers (oracle restatement pinned to oracle/_ref/ref_pack) load
    into the native layout bit-identically."""
    rng = np.random.default_rng(w_bits * 3 + a_bits)
    N, K = 64, 256
    if a_bits < 16:
        hi = (1 << (w_bits - 1)) - 1
        q = rng.integers(-hi, hi + 1, (N, K)).astype(np.int32)
        ref_packed = O.pack_wxax(q, w_bits)
        want = O.pack_wxax_native(q, w_bits)
    else:
        hi = (1 << (w_bits - 1)) - 1
        q = rng.integers(-hi, hi + 1, (N, K)).astype(np.int32) if sym else rng.integers(0, 1 << w_bits, (N, K)).astype(np.int32)
        ref_packed = O.pack_weightonly_ref(q, w_bits, sym)
        want = O.pack_weightonly_native(q, w_bits, sym)
    got = H.repack_from_ref(dev(ref_packed), N, K, w_bits, a_bits, sym)
    torch.cuda.synchronize()
    assert np.array_equal(got.cpu().numpy(), want.reshape(N, -1))


@pytest.mark.parametrize("n,rows,E", [(1408, 1000, 1), (5632, 333, 1), (2560, 517, 1), (1408, 2500, 5), (2816, 130, 3), (10240, 40, 1),
                                      (20480, 700, 1), (14336, 450, 3), (7168, 500, 3), (10240, 333, 5)])
def test_hadamard_pipelined_kernel_bit_identical_to_phase_kernel(H, n, rows, E, monkeypatch):
    """hadamard_pipe_kernel (warp-specialised: cp.async loaders / MMA issuer / TMEM drain + quantize from registers) against
    hadamard_umma_kernel (MXB_HAD_PIPE=0: one phase at a time): the same arithmetic, so identical bits — int8 codes, scales
    and fp16 pass-through rows, with mixed per-expert bit-widths (8 / 4 / 16), ragged row counts (partial last batch), row
    gather, and both replication schemes in use (K = 176: 1408, 5632, 2816; K = 40: 2560).
    Long rows (m = 256: 10240 = 40 x 256, 7168 = 28 x 256; m = 512 as two batches per row: 20480 = 40 x 512, 14336 = 28 x 512;
    several rows per CTA so every barrier changes phase) run on the pipeline too; the other side is then hadamard_wide_kernel
    (one CTA per row, mma.sync): the same butterfly order and fp16 staging, so the same bits are expected here as well."""
    rng = np.random.default_rng(n + rows)
    x = rand16(rng, rows + 7, n, scale=2.0)
    x[3, 7] = 40.0
    src = rng.permutation(rows + 7)[:rows].astype(np.int32)
    if E == 1:
        offs, actq = None, [(8, -1)]
    else:
        cuts = np.sort(rng.integers(0, rows, E - 1))
        offs = dev(np.concatenate([[0], cuts, [rows]]).astype(np.int32))
        actq = [((8, 4, 16)[e % 3], -1) for e in range(E)]

    def run():
        return H.quant_act(dev(x), dev(src), rows, offs, actq, hadamard=True)
    q8, f16, sc = run()
    torch.cuda.synchronize()
    monkeypatch.setenv("MXB_HAD_PIPE", "0")
    q8b, f16b, scb = run()
    torch.cuda.synchronize()
    monkeypatch.delenv("MXB_HAD_PIPE")
    o = np.concatenate([[0], cuts, [rows]]) if E > 1 else np.array([0, rows])
    for e in range(E):
        r0, r1 = int(o[e]), int(o[e + 1])
        if actq[e][0] >= 16:
            assert torch.equal(f16[r0:r1].view(torch.int16), f16b[r0:r1].view(torch.int16))
        else:
            assert torch.equal(q8[r0:r1], q8b[r0:r1])
            assert torch.equal(sc[r0:r1].view(torch.int16), scb[r0:r1].view(torch.int16))
    # and against the oracle, like the phase kernel is: dequantized rows within one quantization step
    h = O.hadamard_f16(x[src])
    e8 = [e for e in range(E) if actq[e][0] == 8]
    if e8:
        r0, r1 = int(o[e8[0]]), int(o[e8[0] + 1])
        if r1 > r0:
            rq, rs = O.quant_act(h[r0:r1], 8, -1)
            deq = q8[r0:r1].cpu().numpy().astype(np.float32) * sc[r0:r1].cpu().numpy().astype(np.float32)
            assert np.abs(deq - rq.astype(np.float32) * rs.astype(np.float32)).max() <= 1.01 * rs.astype(np.float32).max()
            assert np.mean(q8[r0:r1].cpu().numpy().astype(np.int32) != rq) < 0.02


@pytest.mark.parametrize("n,rows", [(2560, 300), (10240, 200), (14336, 180)])
def test_hadamard_kronecker_factor_is_the_same_transform(H, n, rows, monkeypatch):
    """MXB_HAD_KRON (opt-in): the top butterfly stages of the length-m transform folded into the tensor-core factor,
    hadK (x) H_A (2560 = 40 x 64 read as 160 x 16).  The same linear map, but the fp16 staging comes log2(A) stages earlier
    than the reference's (hadamard_utils.py:96-105), so values may differ from the oracle by an fp16 ulp of the intermediate:
    5e-3 of the row's absmax here, against bit-level agreement of the default path with the oracle's staging (tests above).
    This pins that the opt-in path computes the right transform — a wrong sign in H_A would be off by O(1)."""
    rng = np.random.default_rng(n)
    x = rand16(rng, rows, n, scale=2.0)
    monkeypatch.setenv("MXB_HAD_KRON", "16")
    _, f16, _ = H.quant_act(dev(x), None, rows, None, [(16, -1)], hadamard=True)
    q8, _, sc = H.quant_act(dev(x), None, rows, None, [(8, -1)], hadamard=True)
    torch.cuda.synchronize()
    monkeypatch.delenv("MXB_HAD_KRON")
    ref = O.hadamard_f16(x).astype(np.float32)
    got = f16.cpu().numpy().astype(np.float32)
    assert np.abs(got - ref).max() <= 5e-3 * np.abs(ref).max()
    deq = q8.cpu().numpy().astype(np.float32) * sc.cpu().numpy().astype(np.float32).reshape(rows, 1)
    assert np.abs(deq - ref).max() <= 1.1 * sc.cpu().numpy().astype(np.float32).max()
