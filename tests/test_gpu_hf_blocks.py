"""MxMoESparseBlock against the HF sparse-MoE blocks the reference instruments — transformers' own
Qwen2MoeSparseMoeBlock and MixtralSparseMoeBlock, run on the CPU in fp16 like the reference runs its models (linear
outputs, silu * up and the fake-quantized operands are fp16 tensors; an fp32 run of the same block lands one code flip away
on ~6 % of the 8-bit activations, which alone is 2e-3 of the output range here) — prepared the way the reference prepares
them for its accuracy runs:

  * weights fake-quantized offline                Quantizer.fake_quant, mxmoe/quant/quant.py:104-106, :460-461
  * every expert Linear's input fake-quantized    mlp_inp_quant_hook, quant.py:126-130 (per token, the expert's a_bits,
    per token on the fly                          in the model's fp16 arithmetic)
  * down projection rotated                       rotation.py:403-421: W_down <- W_down Q offline, input <- input Q online

against REAL w8a8 / w4a4 experts here (int8 tensor cores over the packed grids).  transformers 5.x keeps the experts as
fused 3-D parameters and calls torch.nn.functional.linear on slices of them, so there is no nn.Linear to hook: the test
patches F.linear for the duration of the HF forward and recognises expert operands by storage address.

Tolerances.  fp16 experts: 4e-3 of the output range (tests/test_gpu_moe_block.py).  Quantized experts: the HF path and the
kernel path are two different roundings of the same model even inside the reference — its CUDA epilogue multiplies the
int32 accumulator by the two scales with fp16 roundings (mm_tile.cuh:445-496, restated by oracle.gemm_wxax), the fake-quant
path rounds an fp32 sum once — and a 5e-4 difference in h ahead of the 8-bit quantizer flips ~4 % of the codes by a full
step, which is as large as the quantization noise itself.  Measured on the CPU between oracle.moe_layer (what the GPU path
is held to bit-for-bit / within 2e-3 in tests/test_gpu_layer*.py) and this HF arm: relative Frobenius error 1.0-1.2 %
(w8a8), 2.1 % (w4a4 + w8a8), max 1.4 % / 3.3 % of the range.  The bars here are 2 % / 4 % Frobenius and 4 % / 8 % max: a
wrong routing weight, expert order, gate/up split, shared-expert gate or rotation is off by O(1).  Tokens whose
top-k set differs between the two routers (logits in fp16 on the GPU, fp16 on the CPU: near ties) are left out and must
stay below 5 %."""
import numpy as np
import pytest
import torch

from oracle import cpu_path as CP
from oracle import oracle as O

pytestmark = pytest.mark.gpu
transformers = pytest.importorskip("transformers")


def lin(w, a):
    return {"w_bits": w, "w_gsize": -1, "w_sym": True, "w_clip": [1.0, 1.0], "a_bits": a, "a_gsize": -1, "a_sym": True, "a_clip": [1.0, 1.0]}


def _bits(kind, e):
    if kind == "fp16":
        return 16
    return (8, 4)[e % 2] if kind == "mixed" else 8


class PatchedLinear:
    """F.linear with the reference's activation hooks for registered weights: {data_ptr: (a_bits, rotate)}"""

    def __init__(self):
        self.reg = {}
        self.orig = torch.nn.functional.linear
        self.hits = 0

    def add(self, w, a_bits, rotate=False):
        self.reg[w.data_ptr()] = (a_bits, rotate)

    def __call__(self, x, w, b=None):
        ent = self.reg.get(w.data_ptr())
        if ent is not None:
            self.hits += 1
            a_bits, rotate = ent
            if rotate:
                # online Hadamard hook with the reference's fp16 staging between the length-m transform and hadK @ (hadamard_utils.py:96-105)
                x = torch.from_numpy(np.asarray(O.hadamard_f16(x.half().numpy()), dtype=np.float16)).to(x.dtype)
            if a_bits < 16:
                x = CP.fake_quant_t(x, a_bits, -1, True)
        return self.orig(x, w, b)


def _prep_w(W, bits, rotate=False):
    """fp16 weight -> (rotated) -> fake-quantized, an fp16 tensor again (the reference's quantized checkpoints are fp16)"""
    Wn = W.half().numpy()
    if rotate:
        Wn = O.hadamard_f16(Wn)
    if bits < 16:
        Wn = O.fake_quant(Wn, bits, -1, True)
    return torch.from_numpy(np.asarray(Wn, dtype=np.float16))


def _run(model, kind, hadamard, monkeypatch):
    from mxmoe_b200 import ops
    from mxmoe_b200.config import qmodel_cfg_from_dict
    from mxmoe_b200.moe_block import MxMoESparseBlock
    qwen = model == "qwen2_moe"
    E, k, H, I, ns, B, S = (8, 4, 256, 1408, 2, 2, 61) if qwen else (8, 2, 256, 448, 0, 2, 61)  # 1408 = 176 x 8, 2816 = 176 x 16, 448 = 28 x 16
    g = torch.Generator().manual_seed(11 + len(kind) + hadamard)
    r = lambda *s, sc=1.0: ((torch.rand(*s, generator=g) * 2 - 1) * sc).half().float()
    experts = [(r(I, H, sc=0.1), r(I, H, sc=0.1), r(H, I, sc=0.1)) for _ in range(E)]
    shared = (r(ns * I, H, sc=0.1), r(ns * I, H, sc=0.1), r(H, ns * I, sc=0.1)) if ns else None
    gate_w, sgate_w = r(E, H, sc=0.3), r(1, H, sc=0.2)
    x = r(B, S, H)
    ex = {str(e): {n: lin(_bits(kind, e), _bits(kind, e)) for n in ("gate", "up", "down")} for e in range(E + (1 if ns else 0))}
    ql = qmodel_cfg_from_dict({"0": {"experts": ex}}).layers["0"]
    blk = MxMoESparseBlock.from_dense_weights(ops.Handle(0), ql, B * S, gate_w, experts, shared, sgate_w if qwen else None, top_k=k,
                                              norm_topk_prob=not qwen, hadamard_down=hadamard)
    out, logits = blk(x.half().cuda())
    torch.cuda.synchronize()
    got = out.reshape(-1, H).float().cpu()

    # ---- the HF block, CPU fp16
    if qwen:
        from transformers.models.qwen2_moe import modeling_qwen2_moe as M
        cfg = M.Qwen2MoeConfig(hidden_size=H, moe_intermediate_size=I, shared_expert_intermediate_size=ns * I, num_experts=E,
                               num_experts_per_tok=k, norm_topk_prob=False, hidden_act="silu")
        hf = M.Qwen2MoeSparseMoeBlock(cfg)
    else:
        from transformers.models.mixtral import modeling_mixtral as M
        cfg = M.MixtralConfig(hidden_size=H, intermediate_size=I, num_local_experts=E, num_experts_per_tok=k, hidden_act="silu")
        hf = M.MixtralSparseMoeBlock(cfg)
    hf = hf.half().eval()
    pl = PatchedLinear()
    with torch.no_grad():
        hf.gate.weight.copy_(gate_w)
        for e, (gw, uw, dw) in enumerate(experts):
            b = _bits(kind, e)
            hf.experts.gate_up_proj[e].copy_(_prep_w(torch.cat([gw, uw], 0), b))
            hf.experts.down_proj[e].copy_(_prep_w(dw, b, hadamard))
            pl.add(hf.experts.gate_up_proj[e], b)
            pl.add(hf.experts.down_proj[e], b, hadamard)
        if qwen:
            b = _bits(kind, E)
            hf.shared_expert.gate_proj.weight.copy_(_prep_w(shared[0], b))
            hf.shared_expert.up_proj.weight.copy_(_prep_w(shared[1], b))
            hf.shared_expert.down_proj.weight.copy_(_prep_w(shared[2], b, hadamard))
            hf.shared_expert_gate.weight.copy_(sgate_w)
            pl.add(hf.shared_expert.gate_proj.weight, b)
            pl.add(hf.shared_expert.up_proj.weight, b)
            pl.add(hf.shared_expert.down_proj.weight, b, hadamard)
        monkeypatch.setattr(torch.nn.functional, "linear", pl)
        res = hf(x.half())
        monkeypatch.undo()
        ref = (res[0] if isinstance(res, tuple) else res).reshape(-1, H).float()
        # routing of the two sides
        w = torch.softmax(torch.nn.functional.linear(x.reshape(-1, H).half(), gate_w.half()).float(), dim=-1)
        ids_hf = torch.topk(w, k, dim=-1).indices.sort(-1).values
    assert pl.hits >= 2 * E  # the hooks really ran
    w_us = torch.softmax(logits.float().cpu(), dim=-1)
    ids_us = torch.topk(w_us, k, dim=-1).indices.sort(-1).values
    same = (ids_hf == ids_us).all(-1)
    assert same.float().mean() >= 0.95
    return got[same], ref[same]


@pytest.mark.parametrize("model", ["qwen2_moe", "mixtral"])
def test_fp16_block_matches_hf_block(model, monkeypatch):
    got, ref = _run(model, "fp16", False, monkeypatch)
    assert torch.isfinite(got).all()
    assert (got - ref).abs().max() <= 4e-3 * ref.abs().max()


@pytest.mark.parametrize("model,kind,hadamard", [("qwen2_moe", "w8a8", False), ("mixtral", "w8a8", False), ("qwen2_moe", "mixed", False),
                                                 ("mixtral", "mixed", False), ("qwen2_moe", "mixed", True), ("mixtral", "w8a8", True)])
def test_quantized_block_matches_fake_quantized_hf_block(model, kind, hadamard, monkeypatch):
    got, ref = _run(model, kind, hadamard, monkeypatch)
    assert torch.isfinite(got).all()
    scale = ref.abs().max()
    frob = (got - ref).norm() / ref.norm()
    mx = (got - ref).abs().max() / scale
    if kind == "mixed":
        assert frob <= 0.04 and mx <= 0.08, (float(frob), float(mx))
    else:
        assert frob <= 0.02 and mx <= 0.04, (float(frob), float(mx))
