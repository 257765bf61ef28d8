"""The MoE-layer hot path as one object: permute -> per-token act-quant -> mixed-precision GroupGEMM (gate_up)
-> silu*mul (+online Hadamard) + quant -> GroupGEMM (down) -> top-k weighted combine.

This is the composition the reference intends with quant_inp_act -> gg_mxmoe_share_fused ->
silu_mul_then_quant -> gg_mxmoe_share_fused (mxmoe/kernels/src/ref_bind.cu:312-757) plus the combine it
lacks (ref_bind.cu:66).  Differences by design: no host synchronisation anywhere (the per-expert token
counts stay on the device and the GEMM kernel reads them through MxbProblem.m_range), problem lists are
marshalled once and re-used through the plan cache, shared experts ride in the same launches.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch

from . import ops
from .config import QExpertConfig, QLayerConfig, QLinearConfig


@dataclass
class ExpertWeights:
    """Packed weights of one expert (index E = the fused shared expert)."""
    gu_packed: torch.Tensor   # gate_up [2I', H] in the native layout of its qconfig
    gu_scale: Optional[torch.Tensor]
    dn_packed: torch.Tensor   # down [H, I']
    dn_scale: Optional[torch.Tensor]
    q_gu: QLinearConfig
    q_dn: QLinearConfig


def _interleave_group_major(s: Optional[torch.Tensor], N: int, asym: bool, inverse: bool = False) -> Optional[torch.Tensor]:
    """Scales in the group-major layout of permute_scale ([K/g, N], asym: [K/g, N, 2] = (scale, zero); stored flat): columns
    [gate_0.. | up_0..] -> (gate_0, up_0, gate_1, up_1, ...) inside every group row, like the weight rows (inverse: back)."""
    if s is None:
        return None
    c = 2 if asym else 1
    assert s.numel() % (N * c) == 0
    v = s.view(-1, N // 2, 2, c) if inverse else s.view(-1, 2, N // 2, c)
    return v.transpose(1, 2).contiguous().view(s.shape)


def _check_tied(q: QExpertConfig):
    g, u = q.gate, q.up
    if (g.w_bits, g.a_bits, g.w_gsize, g.w_sym) != (u.w_bits, u.a_bits, u.w_gsize, u.w_sym):
        raise ValueError("gate and up must share a qconfig (they are one fused GEMM; bits_solver.py:378-379)")


class MoELayer:
    def __init__(self, handle: ops.Handle, E: int, topk: int, H: int, I: int, ns: int, qlayer: QLayerConfig,
                 max_tokens: int, hadamard_down: bool = False, expert_ids: Optional[Sequence[int]] = None, seed: int = 42,
                 weights: Optional[dict] = None, device: Optional[torch.device] = None, weight_scale: float = 1.0,
                 fuse_silu: bool = True, packed: Optional[dict] = None, share: Optional["MoELayer"] = None):
        """expert_ids: the routed experts this rank owns (expert-parallel shard); default all E.
        weights: optional {expert_idx: (W_gate_up fp16 [2I',H], W_down fp16 [H,I'])}, may cover some experts only; the
        others draw U(-1,1) * weight_scale from a generator seeded per expert (the reference bench's distribution,
        test.cu:877-915).
        packed: a stored layer (checkpoint.load_packed(...)[layer]): operands already in their native layouts, copied as
        they are — no fp16 weights and no quantization pass.
        share: another layer of the same stack whose activation / routing buffers this layer re-uses (the layers of a stack
        run one after the other on one stream, so one set of buffers serves all of them; weights stay per layer)."""
        self.h = handle
        self._pool = share._pool if share is not None else {}
        self.E, self.topk, self.H, self.I, self.ns = E, topk, H, I, ns
        self.hadamard_down = hadamard_down
        self.dev = device or torch.device("cuda", handle.device)
        self.local_experts = list(expert_ids) if expert_ids is not None else list(range(E))
        self.EL = len(self.local_experts)
        self.max_tokens = max_tokens
        self.experts: List[ExpertWeights] = []
        ids_all = self.local_experts + ([E] if ns else [])
        # Fuse act_fn(gate)*up into the gate_up GEMM epilogue (MXB_EPI_SILU_MUL) when every problem of the launch is
        # one the CTA-pair kernel covers; otherwise the two-step path (GroupGEMM -> mxb_silu_mul_quant) runs.
        self.fused_silu = fuse_silu and all(self._pair_ok(qlayer.experts[str(e)].gate) for e in ids_all)
        # Decode steps of weight-only 4-bit layers: the streaming kernel gathers the token rows itself (MxbProblem.a_rows)
        # and applies act_fn(gate) * up while staging the down projection's rows (MXB_PRO_SILU_MUL), so the step is
        # permute -> GroupGEMM -> GroupGEMM -> combine.  Eligibility mirrors the kernel choice in api.cu (path 1).
        def _wonly4(q):
            return q.w_bits == 4 and q.a_bits >= 16
        self.stream_fused = (expert_ids is None and max_tokens <= 64 and max_tokens * topk <= 512 and H % 128 == 0 and I % 128 == 0
                             and not hadamard_down
                             and all(_wonly4(qlayer.experts[str(e)].gate) and _wonly4(qlayer.experts[str(e)].down) for e in ids_all))
        # ... and produces act_fn(gate) * up in the gate_up epilogue instead (MXB_EPI_SILU_MUL over interleaved gate/up rows):
        # every element of h is then computed once, not once per CTA that stages the row for the down projection.
        # fuse_silu=False or MXB_STREAM_EPI=0 keeps the staging form (MXB_PRO_SILU_MUL over stacked rows).
        self.stream_epi = self.stream_fused and fuse_silu and os.environ.get("MXB_STREAM_EPI", "1") != "0"
        gen = torch.Generator(device=self.dev)
        ids = self.local_experts + ([E] if ns else [])
        # Stored rows against this layer's row order.  Only the decode form depends on the batch size, so only there the loader
        # adapts: stacked rows keep the staging form; interleaved rows of a decode record are put back in [gate; up] order for a
        # layer that is too large for the streaming kernel.  Everything else is a configuration error.
        relayout = False
        if packed is not None and bool(packed["gate_up_interleaved"]) != bool(self.fused_silu or self.stream_epi):
            if self.stream_epi:
                self.stream_epi = False
            elif fuse_silu and not self.fused_silu and all(_wonly4(qlayer.experts[str(e)].gate) for e in ids_all):
                relayout = True
        if packed is not None and not relayout and bool(packed["gate_up_interleaved"]) != bool(self.fused_silu or self.stream_epi):
            raise ValueError(f"stored gate_up rows are {'interleaved' if packed['gate_up_interleaved'] else 'stacked'} but this layer "
                             f"runs {'with' if self.fused_silu else 'without'} the fused silu*mul epilogue")
        for e in ids:
            qe = qlayer.experts[str(e)]
            _check_tied(qe)
            Ie = I * ns if e == E else I
            if packed is not None:
                er = packed["experts"].get(str(e))
                if er is None:
                    raise ValueError(f"stored layer has no expert {e}")
                if (er["gate_up"], er["down"]) != (qe.gate.qstr(), qe.down.qstr()):
                    raise ValueError(f"expert {e}: stored as {er['gate_up']}/{er['down']}, configured {qe.gate.qstr()}/{qe.down.qstr()}")
                t = {k: (v.to(self.dev).contiguous() if v is not None else None) for k, v in er["tensors"].items()}
                if relayout:
                    t["gu_packed"] = t["gu_packed"].view(Ie, 2, -1).transpose(0, 1).reshape(2 * Ie, -1).contiguous()
                    t["gu_scale"] = _interleave_group_major(t["gu_scale"], 2 * Ie, not qe.gate.w_sym, inverse=True)
                wb_gu, wb_dn = abs(qe.gate.w_bits), abs(qe.down.w_bits)
                if tuple(t["gu_packed"].shape) != (2 * Ie, H * wb_gu // (16 if wb_gu == 16 else 8)) or \
                        tuple(t["dn_packed"].shape) != (H, Ie * wb_dn // (16 if wb_dn == 16 else 8)):
                    raise ValueError(f"expert {e}: stored operand shapes {tuple(t['gu_packed'].shape)}, {tuple(t['dn_packed'].shape)} "
                                     f"do not match H={H}, I={Ie}")
                self.experts.append(ExpertWeights(t["gu_packed"], t["gu_scale"], t["dn_packed"], t["dn_scale"], qe.gate, qe.down))
                continue
            if weights is not None and e in weights:
                Wgu, Wdn = weights[e]
                Wgu, Wdn = Wgu.to(self.dev), Wdn.to(self.dev)
            else:
                # one stream per expert: expert e has the same weights whichever rank owns it and however many experts the
                # layer holds (an expert-parallel shard, the single-GPU layer it is checked against, the replicated shared expert)
                gen.manual_seed(seed * 1000003 + e)
                Wgu = torch.rand(2 * Ie, H, generator=gen, device=self.dev, dtype=torch.float32).mul_(2).sub_(1).mul_(weight_scale).half()
                Wdn = torch.rand(H, Ie, generator=gen, device=self.dev, dtype=torch.float32).mul_(2).sub_(1).mul_(weight_scale).half()
            gu_p, gu_s = self._pack(Wgu, qe.gate)
            if self.fused_silu:  # rows (gate_i, up_i) adjacent: the epilogue pairs neighbouring TMEM lanes
                gu_p = handle.interleave_gate_up(gu_p)
                gu_s = handle.interleave_gate_up(gu_s) if gu_s is not None else None
            elif self.stream_epi:  # same row order for the streaming kernel; its scales are group-major [K/g, N(, 2)]
                gu_p = handle.interleave_gate_up(gu_p)
                gu_s = _interleave_group_major(gu_s, 2 * Ie, not qe.gate.w_sym)
            self.experts.append(ExpertWeights(gu_p, gu_s, *self._pack(Wdn, qe.down), qe.gate, qe.down))
            del Wgu, Wdn
        self._alloc(max_tokens)
        self._build_problems()

    # ------------------------------------------------------------------ setup
    @staticmethod
    def _pair_ok(q: QLinearConfig) -> bool:
        if q.w_bits == 16 and q.a_bits == 16:
            return True
        return q.w_bits == q.a_bits and q.w_bits in (4, 8, -8) and q.w_gsize == -1 and q.w_sym  # -8: MXB_BITS_E4M3

    def _pack(self, W: torch.Tensor, q: QLinearConfig):
        if q.w_bits == 16:
            return W.contiguous(), None
        return self.h.quant_pack_weight(W, q.w_bits, q.a_bits, q.w_gsize, q.w_sym)

    def _buf(self, name: str, shape, dtype, zero: bool = False) -> torch.Tensor:
        """A buffer of the stack's pool: layers built with `share=` get the tensor an earlier layer allocated under the same
        (name, shape, dtype), so the problem lists of every layer of a stack point at one set of activation buffers."""
        ov = getattr(self, "_override", None)
        if ov is not None and name in ov:  # a subclass supplies this buffer (expert-parallel receive buffers in peer-mapped memory)
            return ov[name]
        key = (name, tuple(int(v) for v in shape), dtype)
        t = self._pool.get(key)
        if t is None:
            t = (torch.zeros if zero else torch.empty)(*key[1], dtype=dtype, device=self.dev)
            self._pool[key] = t
        return t

    def _alloc(self, T: int):
        H, I, k = self.H, self.I, self.topk
        R = T * k  # routed rows (capacity: every token could land on this rank)
        f16, i8 = torch.float16, torch.int8
        B = self._buf
        qs = [e.q_gu for e in self.experts[:self.EL]] + [e.q_dn for e in self.experts[:self.EL]]
        any_q = any(q.a_bits < 16 for q in qs)
        any_f = any(q.a_bits >= 16 for q in qs)
        any_q_dn = any(e.q_dn.a_bits < 16 for e in self.experts[:self.EL])
        g_gu = any(q.a_bits < 16 and q.a_gsize == 128 for q in [e.q_gu for e in self.experts[:self.EL]])
        g_dn = any(q.a_bits < 16 and q.a_gsize == 128 for q in [e.q_dn for e in self.experts[:self.EL]])
        self.x_q8 = B("x_q8", (R, H), i8) if any_q else None
        self.x_f16 = B("x_f16", (R, H), f16) if any_f else None
        self.x_sc = B("x_sc", (R, H // 128 if g_gu else 1), f16) if any_q else None
        # gate_up output: [R, 2I] (two-step) or already silu(gate)*up [R, I] (fused epilogue)
        self.gu_out = B("gu_out", (R, I if self.fused_silu else 2 * I), f16)
        # fp16-only down experts without a rotation consume the fused output directly
        self.h_direct = self.fused_silu and not any_q_dn and not self.hadamard_down
        self.h_q8 = B("h_q8", (R, I), i8) if any_q else None
        self.h_f16 = (self.gu_out if self.h_direct else B("h_f16", (R, I), f16)) if any_f else None
        self.h_sc = B("h_sc", (R, I // 128 if g_dn else 1), f16) if any_q else None
        self.dn_out = B("dn_out", (R, H), f16)
        if self.ns:
            s = self.experts[-1]
            Is = I * self.ns
            self.sx_q8 = B("sx_q8", (T, H), i8) if s.q_gu.a_bits < 16 else None
            self.sx_sc = B("sx_sc", (T, H // 128 if s.q_gu.a_gsize == 128 else 1), f16) if s.q_gu.a_bits < 16 else None
            self.s_gu = B("s_gu", (T, Is if self.fused_silu else 2 * Is), f16)
            self.sh_direct = self.fused_silu and s.q_dn.a_bits >= 16 and not self.hadamard_down
            # like the reference workload (gen_workload.py:101) the shared down problem takes the gate config
            self.sh_q8 = B("sh_q8", (T, Is), i8) if s.q_dn.a_bits < 16 else None
            self.sh_f16 = (self.s_gu if self.sh_direct else B("sh_f16", (T, Is), f16)) if s.q_dn.a_bits >= 16 else None
            self.sh_sc = B("sh_sc", (T, Is // 128 if s.q_dn.a_gsize == 128 else 1), f16) if s.q_dn.a_bits < 16 else None
            self.s_dn = B("s_dn", (T, H), f16)
        self.offs = B("offs", (self.EL + 1,), torch.int32, zero=True)
        self.src = B("src", (R,), torch.int32, zero=True)      # permute outputs live in fixed buffers: problem lists
        self.slot = B("slot", (T, k), torch.int32, zero=True)  # (a_rows) and CUDA graphs keep their addresses

    def _problem(self, q: QLinearConfig, A_q8, A_f16, sc, B, sb, C, rows, N, K, m_range, flags=0, c_row=None, a_rows=None):
        A = A_f16 if q.a_bits >= 16 else A_q8
        return ops.Problem(A, B, C, rows, N, K, q.w_bits, q.a_bits, q.w_gsize, q.w_sym if q.w_bits < 16 else True,
                           scale_a=sc if q.a_bits < 16 else None, scale_b=sb, m_range=m_range, flags=flags, c_row=c_row,
                           a_rows=a_rows, lds_a=int(sc.shape[1]) if (q.a_bits < 16 and sc is not None and sc.dim() == 2) else 0)

    def _build_problems(self):
        """Marshal the two problem lists once; M of every routed expert is read on the device from self.offs."""
        H, I, k, T = self.H, self.I, self.topk, self.max_tokens
        R = T * k
        gu, dn = [], []
        for i, ew in enumerate(self.experts[:self.EL]):
            mr = self.offs[i:i + 2]
            if self.stream_fused:  # gate_up is built per input buffer (A = x itself, _finalize_problems); down reads [gate | up]
                dn.append(self._problem(ew.q_dn, None, self.gu_out, None, ew.dn_packed, ew.dn_scale, self.dn_out, R, H, I, mr,
                                        flags=0 if self.stream_epi else ops.PRO_SILU_MUL))
                continue
            gu.append(self._problem(ew.q_gu, self.x_q8, self.x_f16, self.x_sc, ew.gu_packed, ew.gu_scale, self.gu_out, R, 2 * I, H, mr,
                                    flags=ops.EPI_SILU_MUL if self.fused_silu else 0))
            dn.append(self._problem(ew.q_dn, self.h_q8, self.h_f16, self.h_sc, ew.dn_packed, ew.dn_scale, self.dn_out, R, H, I, mr))
        self._gu_routed, self._dn_routed = gu, dn
        self._T_built = None

    def _finalize_problems(self, T: int, x: torch.Tensor):
        # x itself is an operand only when the shared expert takes fp16 activations; lists are kept per (T, x) so that
        # callers alternating between input buffers (HostPipeline) do not re-marshal every step
        key = (T, x.data_ptr() if (self.stream_fused or (self.ns and self.experts[-1].q_gu.a_bits >= 16)) else 0)
        if self._T_built == key:
            return
        if not hasattr(self, "_lists"):
            self._lists = {}
        if key in self._lists:
            self._gu_arr, self._dn_arr, self._keep = self._lists[key]
            self._T_built = key
            return
        gu, dn = list(self._gu_routed), list(self._dn_routed)
        if self.stream_fused:
            R = self.max_tokens * self.topk
            epi = ops.EPI_SILU_MUL if self.stream_epi else 0  # gu_out then holds h = act_fn(gate) * up as dense [R, I] rows
            gu = [self._problem(ew.q_gu, None, x, None, ew.gu_packed, ew.gu_scale, self.gu_out, R, 2 * self.I, self.H,
                                self.offs[i:i + 2], a_rows=self.src, flags=epi) for i, ew in enumerate(self.experts[:self.EL])]
        if self.ns and self.stream_fused:
            s = self.experts[-1]
            Is = self.I * self.ns
            gu.append(self._problem(s.q_gu, None, x, None, s.gu_packed, s.gu_scale, self.s_gu, T, 2 * Is, self.H, None,
                                    flags=ops.EPI_SILU_MUL if self.stream_epi else 0))
            dn.append(self._problem(s.q_dn, None, self.s_gu, None, s.dn_packed, s.dn_scale, self.s_dn, T, self.H, Is, None,
                                    flags=0 if self.stream_epi else ops.PRO_SILU_MUL))
        elif self.ns:
            s = self.experts[-1]
            Is = self.I * self.ns
            A_f16 = x if s.q_gu.a_bits >= 16 else None
            gu.append(self._problem(s.q_gu, self.sx_q8, A_f16, self.sx_sc, s.gu_packed, s.gu_scale, self.s_gu, T, 2 * Is, self.H, None,
                                    flags=ops.EPI_SILU_MUL if self.fused_silu else 0))
            dn.append(self._problem(s.q_dn, self.sh_q8, self.sh_f16, self.sh_sc, s.dn_packed, s.dn_scale, self.s_dn, T, self.H, Is, None))
        self._gu_arr = self.h.pack_problems(gu)
        self._dn_arr = self.h.pack_problems(dn)
        self._keep = (gu, dn)
        if len(self._lists) >= 8:
            self._lists.clear()
        self._lists[key] = (self._gu_arr, self._dn_arr, self._keep)
        self._T_built = key

    # ------------------------------------------------------------------ hot path
    def actq_gu(self):
        return [(e.q_gu.a_bits, e.q_gu.a_gsize) for e in self.experts[:self.EL]]

    def actq_dn(self):
        return [(e.q_dn.a_bits, e.q_dn.a_gsize) for e in self.experts[:self.EL]]

    def forward(self, x: torch.Tensor, topk_ids: torch.Tensor, topk_w: torch.Tensor, out: Optional[torch.Tensor] = None,
                shared_scale: Optional[torch.Tensor] = None):
        """x fp16 [T,H]; topk_ids int32 [T,k] in [0,E) (all experts local); topk_w fp32 [T,k] -> fp16 [T,H].
        shared_scale: optional fp32 [T] gate of the shared expert (Qwen: sigmoid(shared_expert_gate(x)))."""
        T = x.shape[0]
        assert T <= self.max_tokens and self.EL == self.E, "use forward_rows for an expert-parallel shard"
        h = self.h
        _, src, slot = h.permute(topk_ids, self.E, self.offs, self.src[:T * self.topk], self.slot[:T])
        self.run_experts(x, src, T * self.topk, T)
        return h.combine(self.dn_out, slot, topk_w, self.s_dn if self.ns else None, out, shared_scale if self.ns else None)

    def capture(self, x: torch.Tensor, topk_ids: torch.Tensor, topk_w: torch.Tensor, out: torch.Tensor):
        """CUDA-graph one step over FIXED buffers and return the graph (`g.replay()` re-runs it on new contents).
        A decode step (a few dozen tokens) is a chain of ~8 short launches: issued one by one it is bound by the host's
        launch rate, as a graph by the kernels.  Nothing in forward() synchronises or depends on host values (row counts
        stay on the device), so the capture is exact; two eager steps first fill the plan and table caches so that no
        upload happens while capturing."""
        side = torch.cuda.Stream(self.dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                self.forward(x, topk_ids, topk_w, out)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize(self.dev)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            self.forward(x, topk_ids, topk_w, out)
        return g

    def run_experts(self, x: torch.Tensor, src_row: Optional[torch.Tensor], n_rows: int, T_shared: int):
        """Rows already grouped by local expert (self.offs filled): quant -> gate_up -> silu*mul(+had)+quant -> down.
        x is gathered through src_row (None: x rows are already in sorted order).  Results land in self.dn_out
        (routed rows) and self.s_dn (shared expert over the first T_shared rows of x)."""
        h = self.h
        self._finalize_problems(T_shared, x)
        if self.stream_fused:  # gather and silu*mul live in the GEMM kernel's row staging
            h.group_gemm(self._gu_arr)
            h.group_gemm(self._dn_arr)
            return
        h.quant_act(x, src_row, n_rows, self.offs, self.actq_gu(), False, self.x_q8, self.x_f16, self.x_sc)
        if self.ns:
            s = self.experts[-1]
            if s.q_gu.a_bits < 16:
                h.quant_act(x, None, T_shared, None, [(s.q_gu.a_bits, s.q_gu.a_gsize)], False, self.sx_q8, None, self.sx_sc)
        h.group_gemm(self._gu_arr)
        if self.fused_silu:
            # the GEMM epilogue already produced silu(gate)*up: (rotate +) quantize it for the down GEMM
            if not self.h_direct:
                h.quant_act(self.gu_out, None, n_rows, self.offs, self.actq_dn(), self.hadamard_down, self.h_q8, self.h_f16, self.h_sc)
            if self.ns and not self.sh_direct:
                s = self.experts[-1]
                h.quant_act(self.s_gu, None, T_shared, None, [(s.q_dn.a_bits, s.q_dn.a_gsize)], self.hadamard_down, self.sh_q8,
                            self.sh_f16, self.sh_sc)
        else:
            h.silu_mul_quant(self.gu_out, n_rows, self.offs, self.actq_dn(), self.hadamard_down, self.h_q8, self.h_f16, self.h_sc)
            if self.ns:
                s = self.experts[-1]
                h.silu_mul_quant(self.s_gu, T_shared, None, [(s.q_dn.a_bits, s.q_dn.a_gsize)], self.hadamard_down, self.sh_q8,
                                 self.sh_f16, self.sh_sc)
        h.group_gemm(self._dn_arr)

    def gemm_ops(self, counts: Sequence[int], T_shared: int) -> int:
        """sum 2*M*N*K over gate_up + down (+ shared) for the given per-expert token counts (test.cu:109-116)."""
        tot = 0
        for m in counts:
            tot += 2 * m * (2 * self.I) * self.H + 2 * m * self.H * self.I
        if self.ns:
            Is = self.I * self.ns
            tot += 2 * T_shared * (2 * Is) * self.H + 2 * T_shared * self.H * Is
        return tot

    # ------------------------------------------------------------------ bench helpers
    def gemm_only(self):
        """Re-issue the two persistent GroupGEMM launches on the buffers of the last forward (kernel timing)."""
        self.h.group_gemm(self._gu_arr)
        self.h.group_gemm(self._dn_arr)

    def last_counts(self):
        o = self.offs.cpu().tolist()
        return [o[i + 1] - o[i] for i in range(self.EL)]

    def launches_per_step(self, T: Optional[int] = None) -> int:
        # permute (2; 1 when the T*k ids fit one block: permute_small_kernel) + act-quant (1 [+1 shared]) + GroupGEMM (1)
        # + silu*mul+quant (1 [+1 shared]) + GroupGEMM (1) + combine (1)
        perm = 1 if (self.max_tokens if T is None else T) * self.topk <= 2048 else 2
        if self.stream_fused:
            return perm + 3  # permute + GroupGEMM (gather fused) + GroupGEMM (silu*mul fused) + combine
        n = 5 + perm + (2 if self.ns else 0)
        if self.fused_silu:
            n -= (1 if self.h_direct else 0) + (1 if self.ns and self.sh_direct else 0)
        return n

    def weight_bytes(self, active=None) -> int:
        """bytes of packed weights + scales; `active[i]` masks routed local experts (the shared expert always counts)"""
        tot = 0
        for i, e in enumerate(self.experts):
            if active is not None and i < self.EL and not active[i]:
                continue
            for t in (e.gu_packed, e.gu_scale, e.dn_packed, e.dn_scale):
                if t is not None:
                    tot += t.numel() * t.element_size()
        return tot


class MoEStack:
    """The MoE blocks of a model's decoder stack, one after the other on one stream (BASELINE configs[4]: "full MoE stack"):
    per-layer weights and qconfigs (the solver assigns bit-widths layer by layer, ref bits_solver.py:201-422), ONE set of
    activation / routing / exchange buffers (`share=`).  In a decoder the attention block and the norms sit between two MoE
    blocks, so every block's input is a fresh normalised hidden state: forward() feeds the same batch `x` to every block
    with that block's own routing (topk_ids[l], topk_w[l]); `out` holds the last block's result when it returns."""

    def __init__(self, layers: Sequence["MoELayer"]):
        assert len(layers) >= 1
        self.layers = list(layers)
        l0 = self.layers[0]
        self.dev, self.H, self.topk, self.E = l0.dev, l0.H, l0.topk, l0.E

    def forward(self, x: torch.Tensor, topk_ids: torch.Tensor, topk_w: torch.Tensor, out: Optional[torch.Tensor] = None, marks=None):
        """topk_ids int32 [L, T, k], topk_w fp32 [L, T, k].  marks: optional list that receives one CUDA event per block end."""
        for i, layer in enumerate(self.layers):
            out = layer.forward(x, topk_ids[i], topk_w[i], out)
            if marks is not None:
                ev = torch.cuda.Event(enable_timing=True)
                ev.record()
                marks.append(ev)
        return out

    def gemm_only(self):
        for layer in self.layers:
            layer.gemm_only()

    def launches_per_step(self) -> int:
        return sum(layer.launches_per_step() for layer in self.layers)


class HostPipeline:
    """Streams HOST-resident batches through a layer: the call a user with pinned host buffers makes.

    Three streams (H2D copies, compute = the caller's current stream, D2H copies) over `depth` device slots, so the
    upload of batch i+1 and the download of batch i-1 run under the kernels of batch i; every batch still pays its own
    H2D and D2H.  `submit` returns at once; `join()` makes the current stream wait for every download issued so far.
    The reference has no such path (its bench generates inputs on the device, test.cu:877-915)."""

    def __init__(self, layer, T: int, depth: int = 2, graph: bool = False):
        """graph=True: every slot's step is captured once (MoELayer.capture) and replayed — for launch-bound decode steps."""
        d = layer.dev
        self.layer = layer
        self.graph = graph
        self.s_in, self.s_out = torch.cuda.Stream(d), torch.cuda.Stream(d)
        self.slots = []
        stack = isinstance(layer, MoEStack)
        rshape = (len(layer.layers), T, layer.topk) if stack else (T, layer.topk)  # a stack takes every block's routing
        self.tdim = 1 if stack else 0
        for _ in range(depth):
            self.slots.append(dict(
                x=torch.empty(T, layer.H, dtype=torch.float16, device=d), ids=torch.empty(rshape, dtype=torch.int32, device=d),
                w=torch.empty(rshape, dtype=torch.float32, device=d), out=torch.empty(T, layer.H, dtype=torch.float16, device=d),
                up=torch.cuda.Event(), done=torch.cuda.Event(), down=torch.cuda.Event(), g=None, T=None))
        self.n = 0

    def submit(self, x_host: torch.Tensor, ids_host: torch.Tensor, w_host: torch.Tensor, out_host: torch.Tensor):
        sl = self.slots[self.n % len(self.slots)]
        self.n += 1
        cur = torch.cuda.current_stream()
        T = x_host.shape[0]
        self.s_in.wait_event(sl["done"])  # the kernels that last read this slot's inputs are finished
        ids_d, w_d = sl["ids"].narrow(self.tdim, 0, T), sl["w"].narrow(self.tdim, 0, T)
        with torch.cuda.stream(self.s_in):
            sl["x"][:T].copy_(x_host, non_blocking=True)
            ids_d.copy_(ids_host, non_blocking=True)
            w_d.copy_(w_host, non_blocking=True)
            sl["up"].record(self.s_in)
        cur.wait_event(sl["up"])
        cur.wait_event(sl["down"])        # the previous download of this slot's output is finished
        if self.graph:
            if sl["g"] is None or sl["T"] != T:
                sl["g"], sl["T"] = self.layer.capture(sl["x"][:T], ids_d, w_d, sl["out"][:T]), T
            sl["g"].replay()
        else:
            self.layer.forward(sl["x"][:T], ids_d, w_d, sl["out"][:T])
        sl["done"].record(cur)
        self.s_out.wait_event(sl["done"])
        with torch.cuda.stream(self.s_out):
            out_host.copy_(sl["out"][:T], non_blocking=True)
            sl["down"].record(self.s_out)

    def join(self):
        cur = torch.cuda.current_stream()
        for sl in self.slots:
            cur.wait_event(sl["down"])
