"""The sparse-MoE block of a decoder layer on top of the GroupGEMM path: router -> top-k -> experts -> weighted sum,
as a torch.nn.Module with the calling convention of the HF blocks the reference instruments
(Qwen2MoeSparseMoeBlock / MixtralSparseMoeBlock / DeepSeek-V2 MoE).

Routing semantics are the ones the reference's tracer records (mxmoe/quant/moe_tracer.py:79-86):
    routing_weights = softmax(router_logits, dim=1, dtype=float32); top-k; optional renormalisation (norm_topk_prob)
and, for Qwen, the shared expert's output is gated per token by sigmoid(shared_expert_gate(x)).  The router is a
[T, H] x [H, E] product plus a softmax over E <= 64 values — left to torch (plumbing, not on the hot path); everything
from the permute on runs through the C-ABI library.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F

from . import ops
from .config import QLayerConfig
from .layer import MoELayer


def route(router_logits: torch.Tensor, top_k: int, norm_topk_prob: bool) -> Tuple[torch.Tensor, torch.Tensor]:
    """moe_tracer.py:79-86.  Returns (topk_ids int32 [T,k], topk_w fp32 [T,k])."""
    w = F.softmax(router_logits, dim=1, dtype=torch.float)
    w, ids = torch.topk(w, top_k, dim=-1)
    if norm_topk_prob:
        w = w / w.sum(dim=-1, keepdim=True)
    return ids.to(torch.int32).contiguous(), w.contiguous()


def rotate_down_weight(handle: ops.Handle, W_down: torch.Tensor) -> torch.Tensor:
    """The offline half of the online down-projection rotation (rotation.py:403-421: apply_exact_had_to_linear(W,
    had_dim=-1, output=False) before the activation hook is installed): every row of W_down [H, I] — the in_features axis
    — goes through the same Hadamard transform the activations get, so that (x Q)(W Q)^T = x W^T.  Runs on the device through
    the library's own transform in fp16 pass-through mode (the reference rotates in float64 and casts back: the difference
    is the fp16 staging of mxb_quant_act, an ulp of the rotated weight)."""
    W = W_down.to(handle_device(handle), torch.float16).contiguous()
    _, rot, _ = handle.quant_act(W, None, W.shape[0], None, [(16, -1)], hadamard=True)
    return rot


def handle_device(handle: ops.Handle) -> torch.device:
    return torch.device("cuda", getattr(handle, "device", 0) or 0)


class MxMoESparseBlock(torch.nn.Module):
    """forward(hidden_states [B, S, H] or [T, H]) -> (out, router_logits), like the HF sparse-MoE blocks.

    gate_weight: router [E, H]; shared_gate_weight: Qwen's shared_expert_gate [1, H] (None: shared expert, if any, is added
    ungated as in DeepSeek-V2).  `weights` / `qlayer` as for MoELayer: {e: (W_gate_up [2I', H], W_down [H, I'])} with index E
    the (fused) shared expert, and the per-expert qconfig of the reference's JSON.  With hadamard_down=True the W_down given
    HERE must already be in the rotated basis (rotate_down_weight); from_dense_weights rotates plain HF weights itself."""

    def __init__(self, handle: ops.Handle, E: int, top_k: int, H: int, I: int, ns: int, qlayer: QLayerConfig, max_tokens: int,
                 gate_weight: torch.Tensor, weights: Optional[Dict[int, Tuple[torch.Tensor, torch.Tensor]]], norm_topk_prob: bool = False,
                 shared_gate_weight: Optional[torch.Tensor] = None, hadamard_down: bool = False, packed: Optional[dict] = None):
        super().__init__()
        self.layer = MoELayer(handle, E, top_k, H, I, ns, qlayer, max_tokens=max_tokens, hadamard_down=hadamard_down, weights=weights,
                              packed=packed)
        dev = self.layer.dev
        self.top_k, self.norm_topk_prob = top_k, norm_topk_prob
        self.register_buffer("gate_weight", gate_weight.to(dev, torch.float16))
        self.register_buffer("shared_gate_weight", None if shared_gate_weight is None else shared_gate_weight.to(dev, torch.float16))

    @classmethod
    def from_dense_weights(cls, handle, qlayer, max_tokens, gate_weight, expert_weights, shared_weights=None, shared_gate_weight=None,
                           top_k=2, norm_topk_prob=False, hadamard_down=False, weights_prerotated=False):
        """expert_weights: list of (gate_proj [I,H], up_proj [I,H], down_proj [H,I]) per routed expert — the parameters of an
        HF expert MLP; shared_weights: the same triple for the shared expert (intermediate ns*I).  gate and up are stacked
        into the fused gate_up operand the GroupGEMM takes (gen_workload.py:70-76: N = 2I).  hadamard_down=True: every
        down_proj is rotated first (rotate_down_weight) unless weights_prerotated says the caller already did — a
        checkpoint written by the reference's rotation pass holds rotated weights."""
        if hadamard_down and not weights_prerotated:
            expert_weights = [(g, u, rotate_down_weight(handle, d)) for g, u, d in expert_weights]
            if shared_weights is not None:
                shared_weights = (shared_weights[0], shared_weights[1], rotate_down_weight(handle, shared_weights[2]))
        E = len(expert_weights)
        H = expert_weights[0][2].shape[0]
        I = expert_weights[0][0].shape[0]
        weights = {e: (torch.cat([g, u], 0).half(), d.half()) for e, (g, u, d) in enumerate(expert_weights)}
        ns = 0
        if shared_weights is not None:
            g, u, d = shared_weights
            assert g.shape[0] % I == 0, "the shared expert's intermediate size must be a multiple of the routed one"
            ns = g.shape[0] // I
            weights[E] = (torch.cat([g, u], 0).half(), d.half())
        return cls(handle, E, top_k, H, I, ns, qlayer, max_tokens, gate_weight, weights, norm_topk_prob, shared_gate_weight, hadamard_down)

    @classmethod
    def from_packed(cls, handle, record: dict, max_tokens: int, gate_weight, shared_gate_weight=None, norm_topk_prob=False):
        """record: one layer of checkpoint.load_packed(...) — experts start from their stored native-layout operands."""
        from .checkpoint import qlayer_of
        return cls(handle, record["E"], record["top_k"], record["H"], record["I"], record["ns"], qlayer_of(record), max_tokens, gate_weight,
                   None, norm_topk_prob, shared_gate_weight, record["hadamard_down"], packed=record)

    def forward(self, hidden_states: torch.Tensor):
        shape = hidden_states.shape
        x = hidden_states.reshape(-1, shape[-1]).to(torch.float16).contiguous()
        router_logits = F.linear(x, self.gate_weight)
        ids, w = route(router_logits, self.top_k, self.norm_topk_prob)
        scale = None
        if self.shared_gate_weight is not None and self.layer.ns:
            scale = torch.sigmoid(F.linear(x, self.shared_gate_weight).float()).reshape(-1).contiguous()
        out = self.layer.forward(x, ids, w, shared_scale=scale)
        return out.reshape(shape), router_logits
