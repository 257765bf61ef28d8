"""Expert-parallel MoE layer over the GPUs of one node (one process per GPU, torch.distributed / NCCL).

The reference has no multi-GPU path at all (no NCCL/MPI anywhere); this is the partitioning
BASELINE.json's north_star asks for: rank r owns a contiguous slice of the routed experts (weights never
move), every rank brings its own tokens, and the only exchanges are the token DISPATCH (rows already
quantized to int8 containers + fp16 scales by the sender when the experts allow it: half the NVLink bytes
of fp16 rows) and the COMBINE of the fp16 down-projection rows, both `all_to_all_single` with per-peer split
sizes.  Shared experts are replicated and run on the local tokens in the same GroupGEMM launches.

Received rows arrive grouped by (source rank, local expert).  They are NOT regrouped: every (source, expert)
segment becomes its own GroupGEMM problem pointing at the same expert weights, with its row range read on the
device (MxbProblem.m_range), so dispatch -> GEMM needs no gather pass.

The one host synchronisation per step is the read-back of the [W, E] count matrix needed for the split sizes.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.distributed as dist

from . import ops
from .config import QLayerConfig
from .layer import MoELayer


def expert_bounds(E: int, W: int) -> List[int]:
    """Contiguous expert slices: rank r owns [b[r], b[r+1])."""
    return [r * E // W for r in range(W + 1)]


def ep_plan(counts: np.ndarray, bounds: Sequence[int], rank: int):
    """Split sizes of the dispatch from the all-gathered count matrix counts[src_rank, expert].

    Returns (send_splits[W], recv_splits[W], seg_counts[W*EL]) where seg_counts lists, in arrival order
    (source-major, local-expert-minor), how many rows of each local expert every source sends."""
    W = counts.shape[0]
    lo, hi = bounds[rank], bounds[rank + 1]
    send = [int(counts[rank, bounds[p]:bounds[p + 1]].sum()) for p in range(W)]
    recv = [int(counts[s, lo:hi].sum()) for s in range(W)]
    seg = counts[:, lo:hi].reshape(-1).astype(np.int64)
    return send, recv, seg


class EPMoELayer(MoELayer):
    def __init__(self, handle: ops.Handle, E: int, topk: int, H: int, I: int, ns: int, qlayer: QLayerConfig, max_tokens: int,
                 hadamard_down: bool = False, capacity: float = 2.0, seed: int = 42, weight_scale: float = 1.0,
                 group: Optional[dist.ProcessGroup] = None, weights: Optional[dict] = None):
        self.group = group
        self.W = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.bounds = expert_bounds(E, self.W)
        mine = list(range(self.bounds[self.rank], self.bounds[self.rank + 1]))
        self.T_local = max_tokens
        # receive capacity in rows: `capacity` x the balanced share
        self.cap_rows = int(max_tokens * topk * capacity) + 128
        self.q_all = [(qlayer.experts[str(e)].gate.a_bits, qlayer.experts[str(e)].gate.a_gsize) for e in range(E)]
        self.sender_quant = all(a < 16 for a, _ in self.q_all)
        self.any_g128 = any(g == 128 for a, g in self.q_all if a < 16)
        cap_tokens = (self.cap_rows + topk - 1) // topk
        # every rank draws the weights of ALL experts from the same stream position so that the model is the same
        # regardless of W: seed is offset per expert inside MoELayer by construction order -> use a per-rank seed
        super().__init__(handle, E, topk, H, I, ns, qlayer, max_tokens=cap_tokens, hadamard_down=hadamard_down,
                         expert_ids=mine, seed=seed + 1000 * self.rank, weight_scale=weight_scale, weights=weights)
        d = self.dev
        R = max_tokens * topk
        cols = H // 128 if self.any_g128 else 1
        if self.sender_quant and self.x_sc.shape[1] != cols:  # receive buffer must match the senders' scale width
            self.x_sc = torch.empty(self.x_sc.shape[0], cols, dtype=torch.float16, device=d)
        self.send_q8 = torch.empty(R, H, dtype=torch.int8, device=d) if self.sender_quant else None
        self.send_sc = torch.empty(R, cols, dtype=torch.float16, device=d) if self.sender_quant else None
        self.send_f16 = None if self.sender_quant else torch.empty(R, H, dtype=torch.float16, device=d)
        self.recv_f16 = None if self.sender_quant else torch.empty(self.cap_rows, H, dtype=torch.float16, device=d)
        self.back = torch.empty(R, H, dtype=torch.float16, device=d)
        self.seg_offs = torch.zeros(self.W * self.EL + 1, dtype=torch.int32, device=d)
        self.counts_all = torch.zeros(self.W, E, dtype=torch.int32, device=d)
        self._last_seg = None
        self._build_segment_problems()
        # MXB_EP_TIMING=1: CUDA events at the phase boundaries of forward(), read with timing_report()
        import os
        self._timing = os.environ.get("MXB_EP_TIMING", "0") == "1"
        self._marks = []

    def _mark(self, name: str):
        if self._timing:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self._marks.append((name, ev))

    def timing_report(self):
        """mean milliseconds per phase over the forwards recorded so far (synchronises)"""
        torch.cuda.synchronize()
        acc, cnt = {}, {}
        for (n0, e0), (n1, e1) in zip(self._marks[:-1], self._marks[1:]):
            if n1 == "start":
                continue
            acc[n1] = acc.get(n1, 0.0) + e0.elapsed_time(e1)
            cnt[n1] = cnt.get(n1, 0) + 1
        self._marks = []
        return {k: acc[k] / cnt[k] for k in acc}

    # the parent's per-expert lists are replaced by per-(source, expert) segment lists
    def _build_problems(self):
        self._T_built = None

    def _build_segment_problems(self):
        H, I = self.H, self.I
        R = self.x_q8.shape[0] if self.x_q8 is not None else self.x_f16.shape[0]
        gu, dn = [], []
        for s in range(self.W):
            for i, ew in enumerate(self.experts[:self.EL]):
                mr = self.seg_offs[s * self.EL + i: s * self.EL + i + 2]
                gu.append(self._problem(ew.q_gu, self.x_q8, self.x_f16, self.x_sc, ew.gu_packed, ew.gu_scale, self.gu_out, R, 2 * I, H, mr,
                                        flags=ops.EPI_SILU_MUL if self.fused_silu else 0))
                dn.append(self._problem(ew.q_dn, self.h_q8, self.h_f16, self.h_sc, ew.dn_packed, ew.dn_scale, self.dn_out, R, H, I, mr))
        self._gu_routed, self._dn_routed = gu, dn
        self._seg_actq_gu = [(e.q_gu.a_bits, e.q_gu.a_gsize) for _ in range(self.W) for e in self.experts[:self.EL]]
        self._seg_actq_dn = [(e.q_dn.a_bits, e.q_dn.a_gsize) for _ in range(self.W) for e in self.experts[:self.EL]]

    def _a2a(self, out: torch.Tensor, inp: torch.Tensor, out_splits, in_splits):
        dist.all_to_all_single(out, inp, output_split_sizes=out_splits, input_split_sizes=in_splits, group=self.group)

    def forward(self, x: torch.Tensor, topk_ids: torch.Tensor, topk_w: torch.Tensor, out: Optional[torch.Tensor] = None):
        h, T, k, E, W = self.h, x.shape[0], self.topk, self.E, self.W
        n_send = T * k
        self._mark("start")
        offs, src, slot = h.permute(topk_ids, E)
        my_counts = (offs[1:] - offs[:-1]).contiguous()
        dist.all_gather_into_tensor(self.counts_all.view(-1), my_counts, group=self.group)
        self._mark("permute+allgather")
        counts = self.counts_all.cpu().numpy()  # the one host sync of the step
        self._mark("counts_to_host")
        send, recv, seg = ep_plan(counts, self.bounds, self.rank)
        n_recv = int(sum(recv))
        if n_recv > self.cap_rows:
            raise ops.MxbError(f"rank {self.rank}: {n_recv} routed rows exceed the receive capacity {self.cap_rows}; raise `capacity`")
        self.seg_offs.copy_(torch.from_numpy(np.concatenate([[0], np.cumsum(seg)]).astype(np.int32)), non_blocking=True)
        self._last_seg = seg
        # ---- dispatch
        if self.sender_quant:
            h.quant_act(x, src, n_send, offs, self.q_all, False, self.send_q8, None, self.send_sc)
            self._a2a(self.x_q8[:n_recv], self.send_q8[:n_send], recv, send)
            self._a2a(self.x_sc[:n_recv], self.send_sc[:n_send], recv, send)
        else:
            h.quant_act(x, src, n_send, None, [(16, -1)], False, None, self.send_f16, None)  # gather rows into sorted order
            self._a2a(self.recv_f16[:n_recv], self.send_f16[:n_send], recv, send)
            h.quant_act(self.recv_f16, None, n_recv, self.seg_offs, self._seg_actq_gu, False, self.x_q8, self.x_f16, self.x_sc)
        self._mark("quant+dispatch_a2a")
        # ---- local experts on (source, expert) segments + the replicated shared expert on the local tokens
        self._run_segments(x, n_recv, T)
        self._mark("experts")
        # ---- combine
        self._a2a(self.back[:n_send], self.dn_out[:n_recv], send, recv)
        self._mark("combine_a2a")
        y = h.combine(self.back, slot, topk_w, self.s_dn if self.ns else None, out)
        self._mark("combine")
        return y

    def _run_segments(self, x: torch.Tensor, n_recv: int, T: int):
        h = self.h
        self._finalize_problems(T, x)
        if self.ns:
            s = self.experts[-1]
            if s.q_gu.a_bits < 16:
                h.quant_act(x, None, T, None, [(s.q_gu.a_bits, s.q_gu.a_gsize)], False, self.sx_q8, None, self.sx_sc)
        h.group_gemm(self._gu_arr)
        if self.fused_silu:  # the gate_up epilogue already wrote silu(gate)*up
            if not self.h_direct:
                h.quant_act(self.gu_out, None, n_recv, self.seg_offs, self._seg_actq_dn, self.hadamard_down, self.h_q8, self.h_f16,
                            self.h_sc)
            if self.ns and not self.sh_direct:
                s = self.experts[-1]
                h.quant_act(self.s_gu, None, T, None, [(s.q_dn.a_bits, s.q_dn.a_gsize)], self.hadamard_down, self.sh_q8, self.sh_f16,
                            self.sh_sc)
        else:
            h.silu_mul_quant(self.gu_out, n_recv, self.seg_offs, self._seg_actq_dn, self.hadamard_down, self.h_q8, self.h_f16, self.h_sc)
            if self.ns:
                s = self.experts[-1]
                h.silu_mul_quant(self.s_gu, T, None, [(s.q_dn.a_bits, s.q_dn.a_gsize)], self.hadamard_down, self.sh_q8, self.sh_f16,
                                 self.sh_sc)
        h.group_gemm(self._dn_arr)

    # ------------------------------------------------------------------ bench helpers
    def last_counts(self):
        """rows per local expert processed in the last step (summed over sources)"""
        if self._last_seg is None:
            return [0] * self.EL
        return self._last_seg.reshape(self.W, self.EL).sum(0).tolist()

    def launches_per_step(self) -> int:
        n = 8 + (2 if self.ns else 0) + (0 if self.sender_quant else 1)
        if self.fused_silu:
            n -= (1 if self.h_direct else 0) + (1 if self.ns and self.sh_direct else 0)
        return n
