"""Parity of the sm_100a persistent grouped GEMM (through the C-ABI) against the CPU oracle.

Bars (BASELINE.json north_star): int32 accumulators bit-exact; fp16 outputs within max relative
error 1e-3 (max |out-ref| / max |ref|) of the reference GroupGEMM/dequant path.  For the integer
paths the oracle restates the reference's scale order exactly, so the fp16 outputs are compared
bit-for-bit as well.
"""
import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu

REL_TOL = 1e-3  # north_star: max relative error for fp16 outputs


@pytest.fixture(scope="module", params=["auto", "single_cta"])
def H(request):
    """Every test runs twice: with the kernel chosen per list (the CTA-pair kernel where it is eligible) and with
    the single-CTA kernel forced, so both implementations stay under the same parity bar."""
    from mxmoe_b200 import ops
    h = ops.Handle(0)
    h.set_gemm_path(0 if request.param == "auto" else 1)
    return h


@pytest.fixture(scope="module")
def HP():
    """CTA-pair kernel forced: an ineligible list is an error instead of a silent switch of kernels."""
    from mxmoe_b200 import ops
    h = ops.Handle(0)
    h.set_gemm_path(2)
    return h


def dev(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda()


def rand16(rng, *shape):
    return (rng.random(shape) * 2 - 1).astype(np.float16)  # U(-1,1) like test.cu:877-915


def relerr(out, ref):
    out = out.astype(np.float64)
    ref = ref.astype(np.float64)
    return np.abs(out - ref).max() / max(np.abs(ref).max(), 1e-30)


def make_problem(ops, rng, M, N, K, w_bits, a_bits, gsize, sym, tile_n=0):
    """Quantize with the ORACLE, pack into native layouts with the ORACLE packers -> (Problem, ref_out, ref_acc)."""
    A = rand16(rng, M, K)
    W = rand16(rng, N, K)
    Cbuf = torch.full((max(M, 1), N), float("nan"), dtype=torch.float16, device="cuda")
    if w_bits == 16:
        p = ops.Problem(dev(A), dev(W), Cbuf, M, N, K, tile_n=tile_n)
        return p, O.gemm_fp16(A, W), None
    if a_bits == 16:
        q, s, z = O.quant_weight(W, w_bits, gsize, sym)
        packed = O.pack_weightonly_native(q, w_bits, sym)
        sc = O.permute_scale(s, None if sym else z, sym)
        p = ops.Problem(dev(A), dev(packed), Cbuf, M, N, K, w_bits, 16, gsize, sym, scale_b=dev(sc), tile_n=tile_n)
        return p, O.gemm_wonly(A, q, s, z, gsize, sym), None
    if a_bits == O.BITS_E4M3:  # 8-bit float grids (kind::f8f6f4): fp32 accumulation, no integer accumulator to compare
        va, sa = O.quant_e4m3(A)
        vw, sw = O.quant_e4m3(W)
        p = ops.Problem(dev(O.e4m3_encode(va)), dev(O.e4m3_encode(vw)), Cbuf, M, N, K, O.BITS_E4M3, O.BITS_E4M3, -1, True,
                        scale_a=dev(sa), scale_b=dev(sw.reshape(-1)), tile_n=tile_n)
        return p, O.gemm_e4m3(va, sa, vw, sw), None
    qa, sa = O.quant_act(A, a_bits, gsize)
    qw, sw, _ = O.quant_weight(W, w_bits, gsize, True)
    acc, out = O.gemm_wxax(qa, sa, qw, sw, gsize)
    p = ops.Problem(dev(qa.astype(np.int8)), dev(O.pack_wxax_native(qw, w_bits)), Cbuf, M, N, K, w_bits, a_bits, gsize, True,
                    scale_a=dev(sa), scale_b=dev(np.ascontiguousarray(sw.T)), tile_n=tile_n)
    return p, out, acc


def check(problems, refs, accs, acc_bufs=None):
    torch.cuda.synchronize()
    for i, (p, ref) in enumerate(zip(problems, refs)):
        if p.M == 0:
            continue
        out = p.C[:p.M].cpu().numpy()
        assert not np.isnan(out.astype(np.float32)).any(), f"problem {i}: unwritten outputs"
        if accs[i] is not None:
            if acc_bufs is not None:
                got = acc_bufs[i].cpu().numpy().reshape(accs[i].shape)
                assert np.array_equal(got, accs[i]), f"problem {i}: int32 accumulators differ"
            assert np.array_equal(out.view(np.uint16), ref.view(np.uint16)), f"problem {i}: fp16 output not bit-exact"
        else:
            e = relerr(out, ref)
            assert e <= REL_TOL, f"problem {i}: rel err {e}"


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (128, 256, 256), (200, 384, 320), (37, 64, 2048), (1, 2816, 2048)])
def test_fp16_single(H, M, N, K):
    from mxmoe_b200 import ops
    rng = np.random.default_rng(M * 7 + N)
    p, ref, _ = make_problem(ops, rng, M, N, K, 16, 16, -1, True)
    H.group_gemm([p])
    check([p], [ref], [None])


@pytest.mark.parametrize("tile_n", [32, 64, 128, 256])
def test_fp16_tile_n(H, tile_n):
    from mxmoe_b200 import ops
    rng = np.random.default_rng(tile_n)
    p, ref, _ = make_problem(ops, rng, 300, 512, 192, 16, 16, -1, True, tile_n=tile_n)
    H.group_gemm([p])
    check([p], [ref], [None])


@pytest.mark.parametrize("w_bits,gsize", [(8, -1), (4, -1), (4, 128), (8, 128)])
@pytest.mark.parametrize("M,N,K", [(128, 128, 128), (333, 256, 1408), (64, 2048, 2048)])
def test_wxax_bit_exact(H, w_bits, gsize, M, N, K):
    from mxmoe_b200 import ops
    rng = np.random.default_rng(w_bits * 1000 + M)
    p, ref, acc = make_problem(ops, rng, M, N, K, w_bits, w_bits, gsize, True)
    G = acc.shape[0]
    buf = torch.zeros(G * M * N, dtype=torch.int32, device="cuda")
    H.group_gemm([p], acc_out=[buf])
    check([p], [ref], [acc], [buf])


@pytest.mark.parametrize("w_bits", [4, 8])
@pytest.mark.parametrize("gsize", [-1, 128])
@pytest.mark.parametrize("sym", [True, False])
def test_weight_only(H, w_bits, gsize, sym):
    from mxmoe_b200 import ops
    rng = np.random.default_rng(w_bits + 10 * sym + (gsize & 0xff))
    for (M, N, K) in [(128, 128, 256), (50, 320, 1408), (6, 2816, 2048)]:
        p, ref, _ = make_problem(ops, rng, M, N, K, w_bits, 16, gsize, sym)
        H.group_gemm([p])
        check([p], [ref], [None])


def test_mixed_problem_list_one_launch(H):
    """The reference's point: ONE launch whose problem list mixes precisions (w4a4 + w8a8 + fp16 + w4a16),
    ragged M including empty and single-row problems."""
    from mxmoe_b200 import ops
    rng = np.random.default_rng(5)
    specs = [(310, 256, 512, 4, 4, -1, True), (0, 256, 512, 8, 8, -1, True), (129, 256, 512, 8, 8, -1, True),
             (1, 128, 256, 16, 16, -1, True), (257, 384, 256, 4, 16, 128, False), (77, 128, 384, 4, 4, 128, True),
             (512, 512, 512, 8, 8, -1, True), (5, 64, 64, 16, 16, -1, True), (131, 192, 1408, 4, 16, -1, True)]
    ps, refs, accs = [], [], []
    for (M, N, K, w, a, g, s) in specs:
        p, r, acc = make_problem(ops, rng, M, N, K, w, a, g, s)
        ps.append(p), refs.append(r), accs.append(acc)
    arr = H.pack_problems(ps)
    H.group_gemm(arr)
    H.group_gemm(arr)  # second call hits the plan cache
    check(ps, refs, accs)


def test_device_resident_m_ranges(H):
    """Grouped layout of the MoE layer: one shared A / scale / C buffer, per-expert rows given by DEVICE offsets."""
    from mxmoe_b200 import ops
    rng = np.random.default_rng(9)
    E, N, K = 5, 256, 512
    Ms = [130, 0, 7, 300, 64]
    offs = np.concatenate([[0], np.cumsum(Ms)]).astype(np.int32)
    tot = int(offs[-1])
    A = rand16(rng, tot, K)
    qa, sa = O.quant_act(A, 8, -1)
    d_offs = dev(offs)
    d_A, d_sa = dev(qa.astype(np.int8)), dev(sa)
    Cbuf = torch.full((tot, N), float("nan"), dtype=torch.float16, device="cuda")
    ps, refs = [], []
    keep = []
    for e in range(E):
        W = rand16(rng, N, K)
        qw, sw, _ = O.quant_weight(W, 8, -1, True)
        dB, dsb = dev(O.pack_wxax_native(qw, 8)), dev(np.ascontiguousarray(sw.T))
        keep += [dB, dsb]
        ps.append(ops.Problem(d_A, dB, Cbuf, tot, N, K, 8, 8, -1, True, scale_a=d_sa, scale_b=dsb, m_range=d_offs[e:e + 2]))
        refs.append(O.gemm_wxax(qa[offs[e]:offs[e + 1]], sa[offs[e]:offs[e + 1]], qw, sw, -1)[1])
    H.group_gemm(ps)
    torch.cuda.synchronize()
    out = Cbuf.cpu().numpy()
    for e in range(E):
        assert np.array_equal(out[offs[e]:offs[e + 1]].view(np.uint16), refs[e].view(np.uint16)), f"expert {e}"


def test_unsupported_qconfig_is_an_error(H):
    from mxmoe_b200 import ops
    t = torch.zeros(128, 128, dtype=torch.float16, device="cuda")
    with pytest.raises(ops.MxbError, match="not supported"):
        H.group_gemm([ops.Problem(t, t, t, 128, 128, 128, w_bits=3, a_bits=16, scale_b=t)])
    with pytest.raises(ops.MxbError):
        H.group_gemm([ops.Problem(t, t, t, 128, 128, 100)])  # K not a multiple of 32


def test_full_size_int8_property(H):
    """BASELINE config sizes (Qwen1.5-MoE expert, M=546, N=2816, K=2048): int32 accumulators against an exact
    float64 torch matmul of the same quantized operands on the device (size-independent check)."""
    from mxmoe_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(42)
    M, N, K = 546, 2816, 2048
    for w_bits in (8, 4):
        hi = 127 if w_bits == 8 else 7
        qa = torch.randint(-hi, hi + 1, (M, K), generator=g, device="cuda", dtype=torch.int32)
        qw = torch.randint(-hi, hi + 1, (N, K), generator=g, device="cuda", dtype=torch.int32)
        ref = (qa.double() @ qw.double().T).to(torch.int32)
        if w_bits == 8:
            B = qw.to(torch.int8)
        else:
            B = dev(O.pack_wxax_native(qw.cpu().numpy(), 4))
        sa = torch.ones(M, dtype=torch.float16, device="cuda")
        sb = torch.ones(N, dtype=torch.float16, device="cuda")
        Cb = torch.empty(M, N, dtype=torch.float16, device="cuda")
        acc = torch.zeros(M, N, dtype=torch.int32, device="cuda")
        H.group_gemm([ops.Problem(qa.to(torch.int8), B, Cb, M, N, K, w_bits, w_bits, -1, True, scale_a=sa, scale_b=sb)],
                     acc_out=[acc])
        torch.cuda.synchronize()
        assert torch.equal(acc, ref), f"w{w_bits}: int32 accumulators differ at full size"


def test_many_tiles_per_cta_mixed_kinds(H):
    """Regression: every CTA walks MANY tiles whose kinds alternate (converted w4a4 / plain w8a8 / fp16 / w4a16),
    so the smem ring, the converter barriers and the two TMEM accumulators all wrap several times.
    Checked against exact float64 matmuls on the device (int32 accumulators bit-exact; fp16 within 1e-3)."""
    from mxmoe_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(7)
    rng = np.random.default_rng(7)
    K = 512
    probs, checks, accs = [], [], []
    for i in range(12):
        kind = i % 4
        M, N = 700 + 37 * i, 1024
        Cb = torch.full((M, N), float("nan"), dtype=torch.float16, device="cuda")
        if kind in (0, 1):
            bits = 4 if kind == 0 else 8
            hi = (1 << (bits - 1)) - 1
            qa = torch.randint(-hi, hi + 1, (M, K), generator=g, device="cuda", dtype=torch.int32)
            qw = torch.randint(-hi, hi + 1, (N, K), generator=g, device="cuda", dtype=torch.int32)
            B = qw.to(torch.int8) if bits == 8 else dev(O.pack_wxax_native(qw.cpu().numpy(), 4))
            acc = torch.zeros(M, N, dtype=torch.int32, device="cuda")
            probs.append(ops.Problem(qa.to(torch.int8), B, Cb, M, N, K, bits, bits, -1, True,
                                     scale_a=torch.ones(M, dtype=torch.float16, device="cuda"),
                                     scale_b=torch.ones(N, dtype=torch.float16, device="cuda")))
            accs.append(acc)
            checks.append(("int", (qa.double() @ qw.double().T).to(torch.int32), acc))
        elif kind == 2:
            A = (torch.rand(M, K, generator=g, device="cuda") * 2 - 1).half()
            Wt = (torch.rand(N, K, generator=g, device="cuda") * 2 - 1).half()
            probs.append(ops.Problem(A, Wt, Cb, M, N, K))
            accs.append(None)
            checks.append(("f16", A.double() @ Wt.double().T, Cb))
        else:
            A = (torch.rand(M, K, generator=g, device="cuda") * 2 - 1).half()
            Wn = rand16(rng, N, K)
            q, s, z = O.quant_weight(Wn, 4, 128, False)
            probs.append(ops.Problem(A, dev(O.pack_weightonly_native(q, 4, False)), Cb, M, N, K, 4, 16, 128, False,
                                     scale_b=dev(O.permute_scale(s, z, False))))
            accs.append(None)
            Wd = torch.from_numpy(O.dequant_wonly(q, s, z, 128, False)).cuda()
            checks.append(("f16", A.double() @ Wd.double().T, Cb))
    H.group_gemm(probs, acc_out=accs)
    torch.cuda.synchronize()
    for i, (kind, ref, got) in enumerate(checks):
        if kind == "int":
            assert torch.equal(got, ref), f"problem {i}: int32 accumulators differ"
        else:
            err = (got.double() - ref).abs().max().item() / ref.abs().max().item()
            assert err <= REL_TOL, f"problem {i}: rel err {err}"


@pytest.mark.parametrize("w_bits", [16, 8, 4])
@pytest.mark.parametrize("M", [1, 15, 16, 17, 100, 255, 256, 257, 513, 546])
def test_pair_kernel_ragged_tokens(HP, w_bits, M):
    """CTA-pair kernel (cta_group::2, tokens on the UMMA N axis): token extents around every tiling edge
    (multiples of 16, the 256-token tile, equalised chunks), N with a tail (384 = 256 + 128), K = 1408 (11 slabs,
    so the 6-deep ring wraps at an odd count)."""
    from mxmoe_b200 import ops
    rng = np.random.default_rng(w_bits * 100 + M)
    N, K = 384, 1408
    p, ref, acc = make_problem(ops, rng, M, N, K, w_bits, w_bits, -1, True)
    if acc is None:
        HP.group_gemm([p])
        check([p], [ref], [None])
    else:
        buf = torch.zeros(M * N, dtype=torch.int32, device="cuda")
        HP.group_gemm([p], acc_out=[buf])
        check([p], [ref], [acc], [buf])


def test_pair_kernel_many_tiles_mixed_kinds(HP):
    """Every pair walks many tiles whose kinds alternate (in-place converted w4a4 / w8a8 / fp16) over shared buffers
    with DEVICE-resident row ranges, including empty experts: the ring, the remote converter / epilogue barriers and
    both TMEM accumulators wrap many times.  int32 accumulators bit-exact against float64 matmuls on the device."""
    from mxmoe_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(11)
    E, N, K = 24, 1280, 640
    Ms = [0 if e % 7 == 3 else 40 + 97 * e for e in range(E)]
    offs = np.concatenate([[0], np.cumsum(Ms)]).astype(np.int32)
    tot = int(offs[-1])
    d_offs = dev(offs)
    qa8 = torch.randint(-127, 128, (tot, K), generator=g, device="cuda", dtype=torch.int32)
    qa4 = torch.randint(-7, 8, (tot, K), generator=g, device="cuda", dtype=torch.int32)
    Af = (torch.rand(tot, K, generator=g, device="cuda") * 2 - 1).half()
    A8, A4 = qa8.to(torch.int8), qa4.to(torch.int8)
    ones_m = torch.ones(tot, dtype=torch.float16, device="cuda")
    ones_n = torch.ones(N, dtype=torch.float16, device="cuda")
    Cb = torch.full((tot, N), float("nan"), dtype=torch.float16, device="cuda")
    probs, accs, refs, keep = [], [], [], []
    for e in range(E):
        kind = e % 3
        r = slice(int(offs[e]), int(offs[e + 1]))
        if kind == 2:
            Wt = (torch.rand(N, K, generator=g, device="cuda") * 2 - 1).half()
            keep.append(Wt)
            probs.append(ops.Problem(Af, Wt, Cb, tot, N, K, m_range=d_offs[e:e + 2]))
            accs.append(None)
            refs.append(("f16", Af[r].double() @ Wt.double().T))
            continue
        bits = 4 if kind == 0 else 8
        hi = (1 << (bits - 1)) - 1
        qw = torch.randint(-hi, hi + 1, (N, K), generator=g, device="cuda", dtype=torch.int32)
        B = qw.to(torch.int8) if bits == 8 else dev(O.pack_wxax_native(qw.cpu().numpy(), 4))
        keep.append(B)
        acc = torch.zeros(max(Ms[e], 1), N, dtype=torch.int32, device="cuda")
        probs.append(ops.Problem(A8 if bits == 8 else A4, B, Cb, tot, N, K, bits, bits, -1, True, scale_a=ones_m, scale_b=ones_n,
                                 m_range=d_offs[e:e + 2]))
        accs.append(acc)
        refs.append(("int", ((qa8 if bits == 8 else qa4)[r].double() @ qw.double().T).to(torch.int32)))
    arr = HP.pack_problems(probs)
    for _ in range(3):  # repeated launches reuse the plan; results must not depend on leftover smem / TMEM state
        HP.group_gemm(arr, acc_out=accs)
    torch.cuda.synchronize()
    for e, (kind, ref) in enumerate(refs):
        if Ms[e] == 0:
            continue
        r = slice(int(offs[e]), int(offs[e + 1]))
        if kind == "int":
            assert torch.equal(accs[e][:Ms[e]], ref), f"expert {e}: int32 accumulators differ"
        else:
            err = (Cb[r].double() - ref).abs().max().item() / ref.abs().max().item()
            assert err <= REL_TOL, f"expert {e}: rel err {err}"


def test_pair_kernel_rejects_ineligible_list(HP):
    from mxmoe_b200 import ops
    rng = np.random.default_rng(3)
    p, _, _ = make_problem(ops, rng, 64, 128, 256, 4, 4, 128, True)
    with pytest.raises(ops.MxbError, match="CTA pairs"):
        HP.group_gemm([p])


@pytest.mark.parametrize("w_bits", [16, 8, 4])
@pytest.mark.parametrize("M", [1, 31, 200, 546])
def test_pair_kernel_fused_silu_mul_is_bit_identical_to_two_steps(HP, w_bits, M):
    """MXB_EPI_SILU_MUL (gate_up GEMM writes silu(gate)*up over interleaved weight rows) must equal, bit for bit,
    the reference's two-step pipeline: GroupGEMM -> fp16 [M,2I] -> silu_mul (ref_bind.cu:595-757) in fp16 mode; and
    the two-step result is itself checked against the oracle."""
    from mxmoe_b200 import ops
    rng = np.random.default_rng(w_bits * 31 + M)
    I, K = 192, 640
    p, ref, _ = make_problem(ops, rng, M, 2 * I, K, w_bits, w_bits, -1, True)
    HP.group_gemm([p])
    check([p], [ref], [None if w_bits == 16 else None])
    _, two_step, _ = HP.silu_mul_quant(p.C, M, None, [(16, -1)])
    ref_h = O.silu_mul(ref)
    assert np.abs(two_step.cpu().numpy().astype(np.float32) - ref_h.astype(np.float32)).max() <= 2e-3 * max(1.0, np.abs(ref_h).max())
    Cf = torch.full((M, I), float("nan"), dtype=torch.float16, device="cuda")
    pf = ops.Problem(p.A, HP.interleave_gate_up(p.B), Cf, M, 2 * I, K, p.w_bits, p.a_bits, p.gsize, p.sym, scale_a=p.scale_a,
                     scale_b=None if p.scale_b is None else HP.interleave_gate_up(p.scale_b), flags=ops.EPI_SILU_MUL)
    HP.group_gemm([pf])
    torch.cuda.synchronize()
    assert torch.equal(Cf.view(torch.int16), two_step.view(torch.int16)), "fused silu*mul differs from the two-step path"


def test_fused_silu_mul_needs_the_pair_kernel():
    from mxmoe_b200 import ops
    h = ops.Handle(0)
    h.set_gemm_path(1)
    t = torch.zeros(128, 128, dtype=torch.float16, device="cuda")
    with pytest.raises(ops.MxbError, match="CTA-pair"):
        h.group_gemm([ops.Problem(t, t, t, 128, 128, 128, flags=ops.EPI_SILU_MUL)])


@pytest.mark.parametrize("gsize,sym", [(128, False), (128, True), (-1, True), (-1, False)])
@pytest.mark.parametrize("M,N,K", [(1, 16, 128), (6, 2816, 2048), (8, 528, 1408), (9, 512, 1408), (16, 2048, 1408), (3, 1040, 3584)])
def test_w4a16_streaming_kernel(H, gsize, sym, M, N, K):
    """HBM-streaming weight-only kernel (problems with at most 16 rows): bulk-copied weight slabs, weights on the
    m16 side of mma.sync, every ring/tail case (K = 1408 = 2.75 stages, N not a multiple of 512, 1..16 tokens)."""
    from mxmoe_b200 import ops
    rng = np.random.default_rng(M * 13 + N + K + (gsize & 7) + 2 * sym)
    p, ref, _ = make_problem(ops, rng, M, N, K, 4, 16, gsize, sym)
    H.group_gemm([p])
    check([p], [ref], [None])


@pytest.mark.parametrize("gsize,sym", [(128, False), (128, True), (-1, False)])
@pytest.mark.parametrize("M", [1, 6, 8, 13, 16, 40])
def test_w4a16_streaming_fused_silu_mul_is_bit_identical_to_two_steps(H, gsize, sym, M):
    """MXB_EPI_SILU_MUL on the streaming kernel: interleaved gate/up weight rows and scales, C = act_fn(gate) * up [M, I].
    Must equal, bit for bit, the two-step form (GroupGEMM -> [M, 2I] -> mxb_silu_mul_quant in fp16 mode), whose GEMM is
    checked against the oracle."""
    from mxmoe_b200 import ops
    from mxmoe_b200.layer import _interleave_group_major
    rng = np.random.default_rng(M * 7 + (gsize & 3) + sym)
    I, K = 264, 1408  # 33 slabs of 16 interleaved rows: a partial last round
    p, ref, _ = make_problem(ops, rng, M, 2 * I, K, 4, 16, gsize, sym)
    H.group_gemm([p])
    check([p], [ref], [None])
    _, two_step, _ = H.silu_mul_quant(p.C, M, None, [(16, -1)])
    Cf = torch.full((max(M, 1), I), float("nan"), dtype=torch.float16, device="cuda")
    pf = ops.Problem(p.A, H.interleave_gate_up(p.B), Cf, M, 2 * I, K, 4, 16, gsize, sym,
                     scale_b=_interleave_group_major(p.scale_b, 2 * I, not sym), flags=ops.EPI_SILU_MUL)
    H.group_gemm([pf])
    torch.cuda.synchronize()
    assert torch.equal(Cf[:M].view(torch.int16), two_step[:M].view(torch.int16)), "fused silu*mul differs from the two-step path"


def test_w4a16_streaming_many_experts_device_m(H):
    """Decode-shaped list: 40 experts sharing one activation buffer with DEVICE-resident row ranges (0..16 rows each,
    several empty) -> one streaming launch; every expert against the oracle."""
    from mxmoe_b200 import ops
    rng = np.random.default_rng(77)
    E, N, K = 40, 704, 1024
    Ms = [int(v) for v in rng.integers(0, 17, E)]
    Ms[3] = Ms[17] = 0
    offs = np.concatenate([[0], np.cumsum(Ms)]).astype(np.int32)
    tot = int(offs[-1])
    A = rand16(rng, tot, K)
    d_A, d_offs = dev(A), dev(offs)
    Cb = torch.full((tot, N), float("nan"), dtype=torch.float16, device="cuda")
    ps, refs, keep = [], [], []
    for e in range(E):
        W = rand16(rng, N, K)
        q, s, z = O.quant_weight(W, 4, 128, False)
        dB, dsc = dev(O.pack_weightonly_native(q, 4, False)), dev(O.permute_scale(s, z, False))
        keep += [dB, dsc]
        ps.append(ops.Problem(d_A, dB, Cb, 256, N, K, 4, 16, 128, False, scale_b=dsc, m_range=d_offs[e:e + 2]))
        refs.append(O.gemm_wonly(A[offs[e]:offs[e + 1]], q, s, z, 128, False) if Ms[e] else None)
    H.group_gemm(ps)
    torch.cuda.synchronize()
    out = Cb.cpu().numpy()
    for e in range(E):
        if Ms[e]:
            assert relerr(out[offs[e]:offs[e + 1]], refs[e]) <= REL_TOL, f"expert {e}"


def test_select_tiles_reports_the_kernel_choice(H):
    """Blackwell tile selector (replaces tile_config.py:330-610 / bits_solver.py:425-515): path and tile per problem."""
    from mxmoe_b200 import ops
    t = torch.zeros(16, 16, dtype=torch.float16, device="cuda")
    ps = [ops.Problem(t, t, t, 546, 2816, 2048, 8, 8, -1, True, scale_a=t, scale_b=t),       # CTA pair unless forced off
          ops.Problem(t, t, t, 6, 2816, 2048, 4, 16, 128, False, scale_b=t),                 # decode: streaming kernel
          ops.Problem(t, t, t, 546, 2816, 2048, 4, 4, 128, True, scale_a=t, scale_b=t)]      # group-wise: single-CTA kernel
    ch = H.select_tiles(ps)
    assert ch[1].path == 1 and ch[2].path == 0 and ch[2].tile_n == 64
    assert ch[0].path in (0, 2)
    if ch[0].path == 2:
        assert (ch[0].tile_m, ch[0].tile_n) == (256, 192)  # 546 tokens -> 3 equal tiles of 182 -> 192
    assert all(c.est_us > 0 for c in ch)


@pytest.mark.parametrize("M,N,K", [(128, 256, 128), (300, 512, 2048), (37, 2816, 2048), (546, 2048, 1408)])
def test_w8a8_e4m3(H, M, N, K):
    """`w8a8_g-1_sym_E4M3` (tile_config.py:45): E4M3 codes from the oracle, tcgen05.mma kind::f8f6f4 with fp32
    accumulation, the int path's scale order; both kernels (fixture: auto = CTA pairs, single = one CTA).  The products
    are exact and only the accumulation order differs from the float64 oracle: max relative error <= 1e-3."""
    from mxmoe_b200 import ops
    rng = np.random.default_rng(M + N + K)
    p, ref, _ = make_problem(ops, rng, M, N, K, O.BITS_E4M3, O.BITS_E4M3, -1, True)
    H.group_gemm([p])
    check([p], [ref], [None])


def test_e4m3_mixed_list_and_device_quantizers(H):
    """One launch mixing E4M3, int8 and fp16 problems; the E4M3 operands come from the DEVICE quantizers
    (mxb_quant_act with a_bits = MXB_BITS_E4M3, mxb_quant_pack_weight): codes and scales bit-exact against the oracle."""
    from mxmoe_b200 import ops
    rng = np.random.default_rng(99)
    M, N, K = 200, 512, 1024
    A, W = rand16(rng, M, K), rand16(rng, N, K)
    q8, _, sc = H.quant_act(dev(A), None, M, None, [(ops.BITS_E4M3, -1)])
    wp, ws = H.quant_pack_weight(dev(W), ops.BITS_E4M3, ops.BITS_E4M3, -1, True)
    torch.cuda.synchronize()
    va, sa = O.quant_e4m3(A)
    vw, sw = O.quant_e4m3(W)
    assert np.array_equal(q8.cpu().numpy().view(np.uint8), O.e4m3_encode(va))
    assert np.array_equal(sc.cpu().numpy().view(np.uint16).reshape(-1), sa.view(np.uint16).reshape(-1))
    assert np.array_equal(wp.cpu().numpy(), O.e4m3_encode(vw))
    assert np.array_equal(ws.cpu().numpy().view(np.uint16), sw.view(np.uint16).reshape(-1))
    Cbuf = torch.full((M, N), float("nan"), dtype=torch.float16, device="cuda")
    p8 = ops.Problem(q8, wp, Cbuf, M, N, K, ops.BITS_E4M3, ops.BITS_E4M3, -1, True, scale_a=sc, scale_b=ws)
    pi, refi, _ = make_problem(ops, rng, 100, 256, 512, 8, 8, -1, True)
    pf, reff, _ = make_problem(ops, rng, 64, 128, 256, 16, 16, -1, True)
    H.group_gemm([p8, pi, pf])
    check([p8, pi, pf], [O.gemm_e4m3(va, sa, vw, sw), refi, reff], [None, None, None])


# ----------------------------------------------------------------------------------------------------------------------
# The FuncType shim (mxb_group_gemm_reflayout), driven exactly as INTEGRATION.md section 1's adaptor drives it:
# device `half**` tables, device + host dim3 arrays, device + host 16-byte QParams (registry.cuh:28-39,
# quantize.cuh:14-25, caller test.cu:794-814), NULL stream.
import ctypes as _C


class _Dim3(_C.Structure):
    _fields_ = [("x", _C.c_uint32), ("y", _C.c_uint32), ("z", _C.c_uint32)]


class _QParams(_C.Structure):  # {int2 qbits (x = a_bits, y = w_bits); int gsize; bool sym}
    _fields_ = [("a_bits", _C.c_int32), ("w_bits", _C.c_int32), ("gsize", _C.c_int32), ("sym", _C.c_uint8), ("pad", _C.c_uint8 * 3)]


def _call_shim(h, A, B, sa, sb, Cs, shapes, quants):
    P = len(shapes)
    tab = lambda ts: torch.tensor([t.data_ptr() if t is not None else 0 for t in ts], dtype=torch.int64, device="cuda")
    tA, tB, tsa, tsb, tC = tab(A), tab(B), tab(sa), tab(sb), tab(Cs)
    hs = (_Dim3 * P)(*[_Dim3(*s) for s in shapes])
    hq = (_QParams * P)(*[_QParams(a, w, g, 1 if sym else 0) for (w, a, g, sym) in quants])
    ds = torch.frombuffer(bytearray(bytes(hs)), dtype=torch.uint8).cuda()
    dq = torch.frombuffer(bytearray(bytes(hq)), dtype=torch.uint8).cuda()
    torch.cuda.synchronize()
    with torch.cuda.stream(torch.cuda.default_stream()):
        rc = h.lib.mxb_group_gemm_reflayout(h.h, tA.data_ptr(), tB.data_ptr(), tsa.data_ptr(), tsb.data_ptr(), tC.data_ptr(), tC.data_ptr(),
                                            None, None, None, None, ds.data_ptr(), _C.addressof(hs), dq.data_ptr(), _C.addressof(hq), P, None)
    torch.cuda.synchronize()
    return rc


_SHIM_LIST = [  # (M, N, K, w_bits, a_bits, gsize, sym): fp16, w8a8, an EMPTY problem, w4a4, w4a16 group-wise asym, w8a8 again
    (200, 256, 256, 16, 16, -1, True), (300, 384, 512, 8, 8, -1, True), (0, 256, 512, 4, 4, -1, True),
    (130, 256, 512, 4, 4, -1, True), (77, 256, 256, 4, 16, 128, False), (546, 128, 1024, 8, 8, -1, True)]


@pytest.mark.parametrize("ref_operands", [False, True])
def test_functype_shim_mixed_list(ref_operands):
    """ref_operands=False: operands in the native layouts.  True: A and B exactly as a caller of the reference holds them
    (pack_wxax halves, permute_weight + pack_weightonly words; quantize.cuh:318-475) after mxb_set_ref_operands — the shim
    converts A per call and B once per weight pointer (second call: same result from the cached weights)."""
    from mxmoe_b200 import ops
    h = ops.Handle(0)
    rng = np.random.default_rng(31)
    A, B, sa, sb, Cs, shapes, quants, refs = [], [], [], [], [], [], [], []
    for (M, N, K, wb, ab, g, sym) in _SHIM_LIST:
        X, W = rand16(rng, max(M, 1), K)[:M], rand16(rng, N, K)
        shapes.append((M, N, K))
        quants.append((wb, ab, g, sym))
        Cs.append(torch.full((max(M, 1), N), float("nan"), dtype=torch.float16, device="cuda"))
        if wb == 16:
            A.append(dev(X)); B.append(dev(W)); sa.append(None); sb.append(None)
            refs.append(O.gemm_fp16(X, W))
        elif ab == 16:
            q, s, z = O.quant_weight(W, wb, g, sym)
            A.append(dev(X))
            B.append(dev(O.pack_weightonly_ref(q, wb, sym).view(np.int16)) if ref_operands else dev(O.pack_weightonly_native(q, wb, sym)))
            sa.append(None); sb.append(dev(O.permute_scale(s, None if sym else z, sym)))
            refs.append(O.gemm_wonly(X, q, s, z, g, sym))
        else:
            if M == 0:
                qa, s_a = np.zeros((0, K), np.int32), np.zeros((0, 1), np.float16)
            else:
                qa, s_a = O.quant_act(X, ab, g)
            qw, s_w, _ = O.quant_weight(W, wb, g, True)
            if ref_operands:
                A.append(dev(O.pack_wxax(qa, ab).view(np.int16)) if M else torch.zeros(1, dtype=torch.int16, device="cuda"))
                B.append(dev(O.pack_wxax(qw, wb).view(np.int16)))
            else:
                A.append(dev(qa.astype(np.int8)) if M else torch.zeros(1, dtype=torch.int8, device="cuda"))
                B.append(dev(O.pack_wxax_native(qw, wb)))
            sa.append(dev(s_a.reshape(-1)) if M else torch.zeros(1, dtype=torch.float16, device="cuda"))
            sb.append(dev(np.ascontiguousarray(s_w.T).reshape(-1)))
            refs.append(O.gemm_wxax(qa, s_a, qw, s_w, g)[1] if M else None)
    h._check(h.lib.mxb_set_ref_operands(h.h, (1 if ref_operands else 0)), "mxb_set_ref_operands")
    for call in range(2):
        for c in Cs:
            c.fill_(float("nan"))
        assert _call_shim(h, A, B, sa, sb, Cs, shapes, quants) == 0, h.lib.mxb_last_error().decode()
        for (M, N, K, wb, ab, g, sym), c, r in zip(_SHIM_LIST, Cs, refs):
            if M == 0:
                continue
            got = c.cpu().numpy()
            if wb == 16 or ab == 16:
                assert relerr(got, r) <= REL_TOL
            else:
                assert np.array_equal(got.view(np.uint16), r.view(np.uint16)), f"w{wb}a{ab} call {call}"


def test_functype_shim_fixed_tables_and_errors():
    """MXB_REF_TABLES_FIXED: the pointer tables are fetched once per (table addresses, P).  Unsupported QParams are an error
    code, not an exception (the reference throws std::runtime_error, compose_kernel.py:433)."""
    from mxmoe_b200 import ops
    h = ops.Handle(0)
    rng = np.random.default_rng(32)
    M, N, K = 150, 256, 512
    X, W = rand16(rng, M, K), rand16(rng, N, K)
    qa, s_a = O.quant_act(X, 8, -1)
    qw, s_w, _ = O.quant_weight(W, 8, -1, True)
    ref = O.gemm_wxax(qa, s_a, qw, s_w, -1)[1]
    A, B = [dev(qa.astype(np.int8))], [dev(O.pack_wxax_native(qw, 8))]
    sa, sb = [dev(s_a.reshape(-1))], [dev(np.ascontiguousarray(s_w.T).reshape(-1))]
    Cs = [torch.zeros(M, N, dtype=torch.float16, device="cuda")]
    h._check(h.lib.mxb_set_ref_operands(h.h, 2), "mxb_set_ref_operands")
    for _ in range(3):
        Cs[0].zero_()
        assert _call_shim(h, A, B, sa, sb, Cs, [(M, N, K)], [(8, 8, -1, True)]) == 0
        assert np.array_equal(Cs[0].cpu().numpy().view(np.uint16), ref.view(np.uint16))
    assert _call_shim(h, A, B, sa, sb, Cs, [(M, N, K)], [(3, 8, -1, True)]) == -1  # MXB_EINVAL
    assert b"not supported" in h.lib.mxb_last_error()
    assert h.lib.mxb_set_ref_operands(h.h, 8) == -1
