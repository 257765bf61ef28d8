"""The composed layer at the sizes BASELINE.json names (c2 / c3 / c4 / c5) against the CPU oracle, sub-sampled.

The layer runs the FULL batch (all experts active, the kernels the bench runs: CTA-pair GroupGEMM with the fused
silu*mul epilogue, tcgen05 Hadamard+quant, the streaming w4a16 kernel under a CUDA graph, the fused shared expert);
the oracle — numpy, slow — recomputes a few dozen CHECKED tokens.  Every token's output depends only on its own row
and on the weights of its top-k experts (+ the shared expert), so the checked tokens are routed to a handful of
experts whose fp16 weights the test supplies (the other experts keep the layer's seeded random weights and serve the
other tokens); oracle.moe_layer over the checked tokens alone must then reproduce their rows.
Tolerances are the ones of tests/test_gpu_layer.py, stated per test."""
import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _qt(c):
    return (c.w_bits, c.a_bits, c.w_gsize, c.w_sym if c.w_bits < 16 else True)


def _run(model, layer_idx, q, hadamard, T, n_check, n_sel, seed, graph=False, wscale=None):
    import bench
    from mxmoe_b200 import config as C
    from mxmoe_b200 import ops
    from mxmoe_b200.layer import MoELayer
    s = C.MODEL_SHAPES[model]
    E, k, Hd, I, ns = s["E"], s["topk"], s["H"], s["I"], s["ns"]
    ql = bench.qlayer_for(model, layer_idx, q)
    rng = np.random.default_rng(seed)
    sel = sorted(rng.permutation(E)[:n_sel].tolist())  # experts that serve the checked tokens
    weights = {}
    for e in sel + ([E] if ns else []):
        Ie = I * ns if e == E else I
        weights[e] = (((rng.random((2 * Ie, Hd)) * 2 - 1) * (wscale or 2.0 / np.sqrt(Hd))).astype(np.float16),
                      ((rng.random((Hd, Ie)) * 2 - 1) * (wscale or 2.0 / np.sqrt(Ie))).astype(np.float16))
    x = (rng.random((T, Hd)) * 2 - 1).astype(np.float16)
    ids_t, w_t = bench.routing(model, T, seed=seed, device=None)
    ids, w = ids_t.numpy().copy(), w_t.numpy().copy()
    n_check = min(n_check, T)
    for t in range(n_check):  # checked tokens: k distinct experts out of `sel`
        ids[t] = rng.permutation(sel)[:k]
    h = ops.Handle(0)
    layer = MoELayer(h, E, k, Hd, I, ns, ql, max_tokens=T, hadamard_down=hadamard, weight_scale=0.05,
                     weights={e: (torch.from_numpy(a), torch.from_numpy(b)) for e, (a, b) in weights.items()})
    dx, dids, dw = torch.from_numpy(x).cuda(), torch.from_numpy(ids).cuda(), torch.from_numpy(w).cuda()
    out = torch.empty(T, Hd, dtype=torch.float16, device="cuda")
    if graph:
        g = layer.capture(dx, dids, dw, out)
        out.zero_()
        g.replay()
    else:
        layer.forward(dx, dids, dw, out)
    torch.cuda.synchronize()
    got = out[:n_check].cpu().numpy()
    qc = {e: (_qt(ql.experts[str(e)].gate), _qt(ql.experts[str(e)].down)) for e in sel}
    shared = None
    if ns:
        qs = _qt(ql.experts[str(E)].gate)  # the shared down problem takes the gate config (gen_workload.py:101)
        shared = (weights[E][0], weights[E][1], qs, _qt(ql.experts[str(E)].down))
    ref = O.moe_layer(x[:n_check], ids[:n_check], w[:n_check], weights, qc, E, hadamard, shared)
    return layer, got.astype(np.float64), ref.astype(np.float64)


def _check_quantized(o, r):
    """through quantize -> GEMM -> silu -> (Hadamard ->) quantize -> GEMM a one-step code flip (silu's ex2.approx or the
    fp32 butterflies against float64, one fp16 ulp on the element that sets a row's scale) moves an output by one
    quantization step: 2e-2 of the output range, 99 % of the elements within 2e-3 (tests/test_gpu_layer.py)."""
    assert np.isfinite(o).all()
    scale = np.abs(r).max()
    assert scale > 0
    assert np.abs(o - r).max() <= 2e-2 * scale
    assert np.mean(np.abs(o - r) > 2e-3 * scale) < 0.01


def test_c2_layer_full_size_subsampled():
    """BASELINE configs[1]: Qwen1.5-MoE layer 11, 60 experts top-4 + fused shared expert (N = 11264), mixed w4a4 + w8a8
    (5.0 bits), online Hadamard (176 x 8 routed, 176 x 32 shared), T = 8192."""
    layer, o, r = _run("qwen2_moe", 11, "mixed5", True, T=8192, n_check=48, n_sel=5, seed=21)
    assert layer.fused_silu and not layer.stream_fused
    _check_quantized(o, r)


def test_c1_layer_fp16_full_size_subsampled():
    """BASELINE configs[0]: the same layer in fp16 at T = 512; max relative error 1e-3 of the output range per GEMM,
    2e-3 through the two GEMMs and the silu in between."""
    layer, o, r = _run("qwen2_moe", 11, "fp16", False, T=512, n_check=48, n_sel=5, seed=22)
    assert np.isfinite(o).all()
    assert np.abs(o - r).max() <= 2e-3 * np.abs(r).max()


@pytest.mark.parametrize("T", [1, 8, 64])
def test_c3_decode_layer_full_size(T):
    """BASELINE configs[2]: DeepSeek-V2-Lite, 64 experts top-6 + shared expert, w4a16_g128_asym, decode batches through
    the streaming kernel as a CUDA graph (what bench.py --workload c3 times).  Every token is checked; the routed
    experts are limited to 8 (the other 56 receive no rows: empty problems in the launch)."""
    layer, o, r = _run("ds2", 1, "w4a16_g128_asym", False, T=T, n_check=T, n_sel=8, seed=23 + T, graph=True)
    assert layer.stream_fused
    assert np.isfinite(o).all()
    assert np.abs(o - r).max() <= 2e-3 * np.abs(r).max()


def test_c4_layer_full_size_subsampled():
    """BASELINE configs[3]: Mixtral-8x7B, 8 experts top-2, H = 4096, I = 14336, w8a8_g-1_sym, T = 8192 (one GPU's share of
    the expert-parallel runs)."""
    layer, o, r = _run("mixtral", 0, "w8a8_g-1_sym", False, T=8192, n_check=32, n_sel=2, seed=24)
    assert layer.fused_silu
    _check_quantized(o, r)


def test_c5_layer_full_size_subsampled():
    """BASELINE configs[4], one layer: Qwen2-57B-A14B, 64 experts top-8 + shared expert (I = 20480), mixed w4a4 + w8a8,
    online Hadamard 40 x 64 (routed, tcgen05 form) and 40 x 512 (shared, the long-row kernel), T = 2048."""
    layer, o, r = _run("qwen2_moe_57b", 0, "mixed5", True, T=2048, n_check=32, n_sel=9, seed=25)
    assert layer.fused_silu
    _check_quantized(o, r)


@pytest.mark.parametrize("n,rows", [(14336, 96), (20480, 40), (7168, 64)])
def test_long_row_hadamard_quant_matches_oracle(n, rows):
    """The online Hadamard + per-token quantization alone at the long-row sizes: 14336 = 28 x 512 (Mixtral's down input,
    never run by a layer test since c4 has no rotation), 20480 = 40 x 512, 7168 = 28 x 256.  Same bar as
    tests/test_gpu_moe_ops.py: dequantized rows within one quantization step (fp32 butterflies against float64 can move
    a value across a rounding boundary, or the row's scale by one fp16 ulp), fewer than 2 % of the codes differ."""
    from mxmoe_b200 import ops
    rng = np.random.default_rng(n)
    x = (rng.standard_normal((rows, n)) * 0.5).astype(np.float16)
    h = ops.Handle(0)
    for bits in (8, 4):
        q8, _, sc = h.quant_act(torch.from_numpy(x).cuda(), None, rows, None, [(bits, -1)], hadamard=True)
        torch.cuda.synchronize()
        y = O.hadamard_f16(x)
        qr, sr = O.quant_act(y, bits)
        deq = q8.cpu().numpy().astype(np.float32) * sc.cpu().numpy().astype(np.float32).reshape(rows, 1)
        rdeq = qr.astype(np.float32) * sr.astype(np.float32).reshape(rows, 1)
        assert np.abs(deq - rdeq).max() <= 1.01 * sr.astype(np.float32).max()
        assert np.mean(q8.cpu().numpy().astype(np.int32) != qr) < 0.02
