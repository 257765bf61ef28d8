"""MxMoESparseBlock (router + shared-expert gate + the GroupGEMM path) against a plain torch fp32 restatement of the HF
sparse-MoE block forward (Qwen2-MoE style: softmax -> top-k -> (renorm) -> expert MLPs -> weighted sum + sigmoid-gated
shared expert), fp16 experts so that the only differences are accumulation order and fp16 roundings."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def lin16():
    return {"w_bits": 16, "w_gsize": -1, "w_sym": True, "w_clip": [1.0, 1.0], "a_bits": 16, "a_gsize": -1, "a_sym": True, "a_clip": [1.0, 1.0]}


@pytest.mark.parametrize("norm,shared_gate", [(False, True), (True, False)])
def test_block_matches_torch_reference(norm, shared_gate):
    from mxmoe_b200 import ops
    from mxmoe_b200.config import qmodel_cfg_from_dict
    from mxmoe_b200.moe_block import MxMoESparseBlock
    E, k, H, I, ns, B, S = 8, 2, 256, 128, 2, 3, 37
    g = torch.Generator().manual_seed(7)
    r = lambda *s, sc=1.0: ((torch.rand(*s, generator=g) * 2 - 1) * sc)
    experts = [(r(I, H, sc=0.1), r(I, H, sc=0.1), r(H, I, sc=0.1)) for _ in range(E)]
    shared = (r(ns * I, H, sc=0.1), r(ns * I, H, sc=0.1), r(H, ns * I, sc=0.1))
    gate_w, sgate_w = r(E, H, sc=0.2), r(1, H, sc=0.2)
    ex = {str(e): {"gate": lin16(), "up": lin16(), "down": lin16()} for e in range(E + 1)}
    ql = qmodel_cfg_from_dict({"0": {"experts": ex}}).layers["0"]
    blk = MxMoESparseBlock.from_dense_weights(ops.Handle(0), ql, B * S, gate_w, experts, shared, sgate_w if shared_gate else None,
                                              top_k=k, norm_topk_prob=norm)
    x = r(B, S, H).half().cuda()
    out, logits = blk(x)
    torch.cuda.synchronize()
    assert out.shape == x.shape and logits.shape == (B * S, E)
    # ---- torch fp32 reference over the same fp16-rounded parameters and the block's own router logits
    xf = x.reshape(-1, H).float().cpu()
    h16 = lambda t: t.half().float()
    w = torch.softmax(logits.float().cpu(), dim=1)
    tw, ti = torch.topk(w, k, dim=-1)
    if norm:
        tw = tw / tw.sum(-1, keepdim=True)

    def mlp(xr, gw, uw, dw):
        gu = h16(xr @ h16(torch.cat([gw, uw], 0)).T)
        hmid = h16(h16(torch.nn.functional.silu(gu[:, :gw.shape[0]])) * gu[:, gw.shape[0]:])
        return h16(hmid @ h16(dw).T)
    ref = torch.zeros(B * S, H)
    for e in range(E):
        rows, slots = (ti == e).nonzero(as_tuple=True)
        if len(rows):
            ref[rows] += tw[rows, slots, None] * mlp(xf[rows], *experts[e])
    sh = mlp(xf, *shared)
    if shared_gate:
        sh = torch.sigmoid(xf @ h16(sgate_w).T) * sh
    ref = ref + sh
    got = out.reshape(-1, H).float().cpu()
    scale = ref.abs().max()
    assert (got - ref).abs().max() <= 4e-3 * scale
    assert torch.isfinite(got).all()
