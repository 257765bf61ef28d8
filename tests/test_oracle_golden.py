"""The oracle against fixtures produced by the reference itself (tests/golden/gen_golden.py)."""
import os

import numpy as np
import pytest

from oracle import oracle as O

G = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("bits,gs,sym", [(8, -1, True), (4, -1, True), (4, 128, True), (4, 128, False),
                                         (4, -1, False), (8, 128, False), (8, -1, False), (2, 128, False)])
def test_quant_minmax_matches_reference(bits, gs, sym):
    d = np.load(os.path.join(G, "quant_minmax.npz"))
    key = f"b{bits}_g{gs}_{'sym' if sym else 'asym'}"
    q, s, z = O.quant_minmax(d["x"], bits, gs, sym)
    assert np.array_equal(s, d[key + "_s"])
    assert np.array_equal(z, d[key + "_z"])
    assert np.array_equal(q, d[key + "_q"])
    fq = O.fake_quant(d["x"], bits, gs, sym)
    assert np.array_equal(fq.view(np.uint16), d[key + "_fq"].view(np.uint16))


@pytest.mark.parametrize("n", [64, 2048, 1408, 2560, 1792, 5632, 96])
def test_hadamard_matches_reference(n):
    d = np.load(os.path.join(G, "hadamard.npz"))
    _, k = O.get_hadK(n)
    assert k == int(d[f"k{n}"])
    y = O.hadamard(d[f"x{n}"])
    # the reference divides by a float32 sqrt(n) (hadamard_utils.py:83) -> 1e-7 relative
    np.testing.assert_allclose(y, d[f"y{n}"], rtol=2e-7, atol=1e-9)
    # orthogonality: H H^T = I  -> applying twice with the transpose restores x for K==1
    if k == 1:
        np.testing.assert_allclose(O.hadamard(y), d[f"x{n}"], atol=1e-12)


def test_hadk_selection_order():
    # hadamard_utils.py:39-50: 28 is tested before 40, 40 before 20
    assert O.get_hadK(14336)[1] == 28
    assert O.get_hadK(2560)[1] == 40
    assert O.get_hadK(20480)[1] == 40
    assert O.get_hadK(1408)[1] == 176
    assert O.get_hadK(5632)[1] == 176
    assert O.get_hadK(4096)[1] == 1


@pytest.mark.parametrize("bits", [8, 4, 2])
@pytest.mark.parametrize("sym", [True, False])
def test_pack_weightonly_matches_reference(bits, sym):
    d = np.load(os.path.join(G, "pack.npz"))
    key = f"wonly_b{bits}_{'sym' if sym else 'asym'}"
    q, p = d[key + "_q"], d[key + "_p"]
    assert np.array_equal(O.pack_weightonly_ref(q, bits, sym), p)
    assert np.array_equal(O.unpack_weightonly_ref(p, bits, sym, *q.shape), q)


@pytest.mark.parametrize("bits", [8, 4])
def test_pack_wxax_matches_reference(bits):
    d = np.load(os.path.join(G, "pack.npz"))
    for t in ("a", "w"):
        q, p = d[f"wxax_b{bits}_{t}"], d[f"wxax_b{bits}_p{t}"]
        assert np.array_equal(O.pack_wxax(q, bits), p)
        assert np.array_equal(O.unpack_wxax(p, bits), q)


@pytest.mark.parametrize("sym", [True, False])
def test_permute_scale_matches_reference(sym):
    d = np.load(os.path.join(G, "pack.npz"))
    k = "sym" if sym else "asym"
    s = d[f"scale_{k}_in"]
    out = O.permute_scale(s[..., 0], None if sym else s[..., 1], sym)
    ref = d[f"scale_{k}_out"]
    assert np.array_equal(out.reshape(ref.shape), ref)


def test_gemm_wxax_int_exact_and_order():
    rng = np.random.default_rng(0)
    A = (rng.random((40, 256)) * 2 - 1).astype(np.float16)
    W = (rng.random((48, 256)) * 2 - 1).astype(np.float16)
    for bits, gs in ((8, -1), (4, -1), (4, 128)):
        qa, sa = O.quant_act(A, bits, gs)
        qw, sw, _ = O.quant_weight(W, bits, gs, True)
        acc, out = O.gemm_wxax(qa, sa, qw, sw, gs)
        g = 256 if gs == -1 else gs
        for i in range(256 // g):
            ref = qa[:, i * g:(i + 1) * g].astype(np.int64) @ qw[:, i * g:(i + 1) * g].astype(np.int64).T
            assert np.array_equal(acc[i], ref)
        # close to the un-quantized product for 8 bit
        if bits == 8:
            full = A.astype(np.float64) @ W.astype(np.float64).T
            assert np.abs(out.astype(np.float64) - full).max() < 0.25


def test_permute_and_combine_roundtrip():
    rng = np.random.default_rng(1)
    T, k, E, H = 37, 4, 12, 16
    ids = np.stack([rng.permutation(E)[:k] for _ in range(T)]).astype(np.int32)
    offs, src, slot = O.permute(ids, E)
    assert offs[-1] == T * k
    # rows of expert e are contiguous and keep token order (stable)
    flat = ids.reshape(-1)
    for e in range(E):
        seg = src[offs[e]:offs[e + 1]]
        assert np.all(np.diff(seg) >= 0)
        assert np.all([e in ids[t] for t in seg])
    x = rng.standard_normal((T, H)).astype(np.float16)
    w = np.full((T, k), 0.25, np.float32)
    out = O.combine(x[src], slot, w)
    np.testing.assert_allclose(out.astype(np.float32), x.astype(np.float32), atol=2e-3)
    assert flat[np.argsort(flat, kind="stable")][slot.reshape(-1)].tolist() == flat.tolist()


def _ref_pack(op, N, K, M, qbits, sym, gsize, payload: bytes) -> bytes:
    import subprocess
    exe = os.path.join(os.path.dirname(__file__), "..", "oracle", "_ref", "ref_pack")
    hdr = np.array([op, N, K, M, qbits, int(sym), gsize], np.int32).tobytes()
    return subprocess.run([exe], input=hdr + payload, stdout=subprocess.PIPE, check=True).stdout


@pytest.mark.parametrize("bits,sym", [(4, True), (4, False), (8, True), (2, False)])
def test_packers_against_live_reference_binary(bits, sym):
    """When oracle/_ref/ref_pack (the reference's own packers compiled in place by oracle/Makefile) is present,
    check the restatement on fresh random data as well, not only on the committed goldens."""
    exe = os.path.join(os.path.dirname(__file__), "..", "oracle", "_ref", "ref_pack")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/ref_pack not built (no reference tree on this box)")
    rng = np.random.default_rng(bits * 2 + sym)
    N, K = 64, 96 if bits != 2 else 64
    hi = (1 << (bits - 1)) - 1
    q = (rng.integers(-hi, hi + 1, (N, K)) if sym else rng.integers(0, 1 << bits, (N, K))).astype(np.int32)
    out = np.frombuffer(_ref_pack(0, N, K, 0, bits, sym, -1, q.tobytes()), np.uint16)
    assert np.array_equal(O.pack_weightonly_ref(q, bits, sym).reshape(-1), out)
    if bits in (4, 8) and sym:
        a = rng.integers(-hi, hi + 1, (16, K)).astype(np.int32)
        o = np.frombuffer(_ref_pack(1, N, K, 16, bits, True, -1, a.tobytes() + q.tobytes()), np.uint16)
        pack = 16 // bits
        assert np.array_equal(O.pack_wxax(a, bits).reshape(-1), o[:16 * K // pack])
        assert np.array_equal(O.pack_wxax(q, bits).reshape(-1), o[16 * K // pack:])


def test_e4m3_grid_matches_torch_float8():
    """The oracle's E4M3 rounding (round to nearest even, saturating) and byte encoding against torch's own
    float8_e4m3fn conversion on CPU — the only independent implementation in this image; the reference itself only names
    the kind (tile_config.py:45) and has no quantizer for it."""
    import torch
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.standard_normal(200000) * 100, rng.standard_normal(50000) * 0.01, [0.0, 448.0, -448.0, 500.0, 2 ** -9, 2 ** -10]])
    x = x.astype(np.float32)
    t = torch.from_numpy(x).clamp(-448, 448).to(torch.float8_e4m3fn)
    r = O.e4m3_round(x)
    assert np.array_equal(r, t.float().numpy().astype(np.float64))
    assert np.array_equal(O.e4m3_encode(r), t.view(torch.uint8).numpy())  # including the sign of values that round to zero


def test_e4m3_qstr_roundtrip():
    from mxmoe_b200.config import BITS_E4M3, QLinearConfig
    q = QLinearConfig.from_qstr("w8a8_g-1_sym_E4M3")
    assert (q.w_bits, q.a_bits, q.w_gsize, q.w_sym) == (BITS_E4M3, BITS_E4M3, -1, True)
    assert q.qstr() == "w8a8_g-1_sym_E4M3"
    with pytest.raises(ValueError):
        QLinearConfig.from_qstr("w4a4_g-1_sym_E4M3")


@pytest.mark.parametrize("bits,gsize,sym", [(8, -1, True), (4, -1, True), (4, 128, False), (8, 128, True), (-8, -1, True)])
def test_cpu_arm_fake_quant_is_the_oracles(bits, gsize, sym):
    """oracle/cpu_path.py (the torch code bench.py's CPU legs time) against oracle.fake_quant (pinned above against the
    reference's Quantizer.fake_quant): identical values in fp16 arithmetic, for the integer grids and for the E4M3 kind
    (bits = -8), so the timed baseline computes what the checker computes."""
    import torch
    from oracle import cpu_path as CP
    rng = np.random.default_rng(5 + abs(bits))
    x = (rng.standard_normal((48, 512)) * 0.7).astype(np.float16)
    x[3] = 0  # all-zero row: the E4M3 kind guards scale == 0 like quant_weight (quantize.cuh:256); the int grids give NaN on both sides
    # (the E4M3 kind is compared as the CPU arm runs it, on fp32 tensors: the oracle returns code * scale unrounded)
    xt = torch.from_numpy(x).float() if bits < 0 else torch.from_numpy(x)
    got = CP.fake_quant_t(xt, bits, gsize, sym).numpy().astype(np.float32)
    want = np.asarray(O.fake_quant(x, bits, gsize, sym), dtype=np.float32)
    ok = np.isfinite(want)
    assert np.array_equal(np.isfinite(got), ok)
    assert np.array_equal(got[ok], want[ok])
