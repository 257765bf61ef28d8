"""Packed-expert files through the GPU path: a layer saved in its native layouts and loaded back computes bit-identical
results without touching fp16 weights; the converter re-packs a fake-quantized state dict (reference quant.py:772-794)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def lin(w, a, g=-1, sym=True):
    return {"w_bits": w, "w_gsize": g, "w_sym": sym, "w_clip": [1.0, 1.0], "a_bits": a, "a_gsize": g, "a_sym": True, "a_clip": [1.0, 1.0]}


def _mixed_layer_cfg(E, kinds):
    from mxmoe_b200.config import qmodel_cfg_from_dict
    ex = {}
    for e in range(E + 1):
        c = kinds[e % len(kinds)]
        ex[str(e)] = {"gate": dict(c), "up": dict(c), "down": dict(c)}
    return qmodel_cfg_from_dict({"0": {"experts": ex}}).layers["0"]


@pytest.mark.parametrize("kinds,T", [
    ([lin(4, 4), lin(8, 8)], 300),                      # c2-like mix on the CTA-pair kernel (interleaved gate_up rows)
    ([lin(4, 16, 128, False), lin(8, 8), lin(16, 16)], 200),  # mixed list on the single-CTA kernel (stacked rows)
    ([lin(4, 16, 128, False)], 24),                     # decode step on the streaming kernel
])
def test_saved_layer_reloads_bit_identical(tmp_path, kinds, T):
    from mxmoe_b200 import ops, checkpoint as CK
    from mxmoe_b200.layer import MoELayer
    E, k, H, I, ns = 6, 2, 256, 128, 2
    ql = _mixed_layer_cfg(E, kinds)
    h = ops.Handle(0)
    L = MoELayer(h, E, k, H, I, ns, ql, max_tokens=T, weight_scale=0.1, seed=5)
    p = str(tmp_path / "layer.mxb")
    nbytes = CK.save_packed(p, {7: CK.layer_record(L)}, meta={"note": "unit"})
    # the file is the packed operands, not fp16 weights
    dense = (E + ns) * 3 * H * I * 2
    if all(c["w_bits"] < 16 for c in kinds):
        assert nbytes < dense * 0.6
    rec = CK.load_packed(p, device="cuda")[7]
    assert rec["gate_up_interleaved"] == (L.fused_silu or L.stream_epi)
    assert L.stream_epi == (T == 24)  # the decode step keeps gate/up rows interleaved for the streaming kernel's epilogue
    L2 = CK.layer_from_record(ops.Handle(0), rec, max_tokens=T)
    rng = np.random.default_rng(3)
    x = torch.from_numpy((rng.random((T, H)) * 2 - 1).astype(np.float16)).cuda()
    ids = torch.from_numpy(np.stack([rng.permutation(E)[:k] for _ in range(T)]).astype(np.int32)).cuda()
    w = torch.from_numpy(rng.random((T, k)).astype(np.float32)).cuda()
    y1 = L.forward(x, ids, w).clone()
    y2 = L2.forward(x, ids, w).clone()
    torch.cuda.synchronize()
    assert torch.isfinite(y1.float()).all() and float(y1.float().abs().max()) > 0
    assert torch.equal(y1, y2)
    if L.stream_epi:
        # a decode record (interleaved rows for the streaming kernel's epilogue) also loads into a layer that is too large for
        # that kernel: the loader puts the rows back in [gate; up] order
        Tb = 200
        big = MoELayer(h, E, k, H, I, ns, ql, max_tokens=Tb, weight_scale=0.1, seed=5)
        big2 = CK.layer_from_record(ops.Handle(0), rec, max_tokens=Tb)
        assert not big.stream_fused and not big2.stream_fused
        xb = torch.from_numpy((rng.random((Tb, H)) * 2 - 1).astype(np.float16)).cuda()
        idb = torch.from_numpy(np.stack([rng.permutation(E)[:k] for _ in range(Tb)]).astype(np.int32)).cuda()
        wb = torch.from_numpy(rng.random((Tb, k)).astype(np.float32)).cuda()
        yb, yb2 = big.forward(xb, idb, wb).clone(), big2.forward(xb, idb, wb).clone()
        torch.cuda.synchronize()
        assert torch.equal(yb, yb2)


def test_layout_mismatch_is_refused(tmp_path):
    from mxmoe_b200 import ops, checkpoint as CK
    from mxmoe_b200.layer import MoELayer
    E, k, H, I = 4, 2, 256, 128
    ql = _mixed_layer_cfg(E, [lin(8, 8)])
    h = ops.Handle(0)
    L = MoELayer(h, E, k, H, I, 0, ql, max_tokens=64, weight_scale=0.1)
    rec = CK.layer_record(L)
    assert rec["gate_up_interleaved"]
    with pytest.raises(ValueError, match="interleaved"):
        MoELayer(h, E, k, H, I, 0, ql, max_tokens=64, fuse_silu=False, packed=rec)
    with pytest.raises(ValueError, match="configured"):
        MoELayer(h, E, k, H, I, 0, _mixed_layer_cfg(E, [lin(4, 4)]), max_tokens=64, packed=rec)
    with pytest.raises(ValueError, match="shapes"):
        MoELayer(h, E, k, H, 2 * I, 0, ql, max_tokens=64, packed=rec)


def test_convert_fake_quant_state_dict(tmp_path):
    """Weights that already lie on their quantization grid come back as the same grid points (to one fp16 ulp of the scale)."""
    from mxmoe_b200 import ops, checkpoint as CK
    from oracle import oracle as O
    E, k, H, I, ns, Lidx = 3, 2, 256, 128, 1, 4
    kinds = [lin(4, 16, 128, False), lin(8, 8), lin(4, 4)]
    ql = _mixed_layer_cfg(E, kinds)
    rng = np.random.default_rng(9)
    sd, raw = {}, {}
    for e in range(E + 1):
        c = kinds[e % len(kinds)]
        Ie = I * ns if e == E else I
        for proj, shape in (("gate_proj", (Ie, H)), ("up_proj", (Ie, H)), ("down_proj", (H, Ie))):
            W = ((rng.random(shape) * 2 - 1) * 0.1).astype(np.float16)
            Wq = np.asarray(O.fake_quant(W, c["w_bits"], c["w_gsize"], c["w_sym"])).astype(np.float16)
            name = (f"model.layers.{Lidx}.mlp.experts.{e}.{proj}.weight" if e < E else f"model.layers.{Lidx}.mlp.shared_expert.{proj}.weight")
            sd[name] = torch.from_numpy(Wq)
            raw[(e, proj)] = W
    h = ops.Handle(0)
    rec = CK.convert_fake_quant_state_dict(h, sd, ql, Lidx, E, k, max_tokens=32)
    assert rec["ns"] == ns and rec["H"] == H and rec["I"] == I
    p = str(tmp_path / "conv.mxb")
    CK.save_packed(p, {Lidx: rec})
    L = CK.layer_from_record(h, CK.load_packed(p, device="cuda")[Lidx], max_tokens=32)
    # Re-quantization is idempotent: the layer built from the ORIGINAL fp16 weights with the same qconfig lands on the same
    # grid points, so both layers agree to the one-ulp rounding of the recomputed fp16 scales.
    from mxmoe_b200.layer import MoELayer
    orig = {e: (torch.from_numpy(np.concatenate([raw[(e, "gate_proj")], raw[(e, "up_proj")]], 0)), torch.from_numpy(raw[(e, "down_proj")]))
            for e in range(E + 1)}
    L0 = MoELayer(h, E, k, H, I, ns, ql, max_tokens=32, weights=orig)
    T = 32
    x = torch.from_numpy((rng.random((T, H)) * 2 - 1).astype(np.float16)).cuda()
    ids = torch.from_numpy(np.stack([rng.permutation(E)[:k] for _ in range(T)]).astype(np.int32)).cuda()
    w = torch.from_numpy(rng.random((T, k)).astype(np.float32)).cuda()
    y = L.forward(x, ids, w).float().cpu()
    y0 = L0.forward(x, ids, w).float().cpu()
    rel = float((y - y0).abs().max() / y0.abs().max())
    assert float(y0.abs().max()) > 0 and rel < 5e-3, rel
