"""Expert-parallel MoE layer over the GPUs of one node (one process per GPU; torch.distributed for the rendezvous).

The reference has no multi-GPU path at all (no NCCL/MPI anywhere); this is the partitioning
BASELINE.json's north_star asks for: rank r owns a slice of the routed experts (weights never move), every rank
brings its own tokens, and the only exchanges are the token DISPATCH (rows already quantized to int8 containers +
fp16 scales by the sender when the experts allow it: half the NVLink bytes of fp16 rows) and the COMBINE of the fp16
down-projection rows.  The shared expert is replicated and runs on the local tokens.

Received rows arrive grouped by (source rank, local expert).  They are NOT regrouped: every (source, expert)
segment becomes its own GroupGEMM problem pointing at the same expert weights, with its row range read on the
device (MxbProblem.m_range), so dispatch -> GEMM needs no gather pass.

exchange="push" (the design; default on NCCL process groups): no NCCL and no host read-back on the data path.
Receive buffers, the return buffer and the [W, E] count matrix live in torch symmetric memory (every rank maps every
peer's allocation over NVLink), and the exchange is done by the kernels that produce the data:
  counts     mxb_ep_push_counts writes this rank's counts into every peer's matrix            -> barrier
  plan       mxb_ep_plan derives, ON THE DEVICE, the destination of every expert's rows, the layout of what this rank
             receives (segment offsets = the GEMMs' m_range) and where each segment's results return to
  dispatch   mxb_quant_act_push = gather + per-token quantization whose stores go straight into the owner's buffers
                                                                                               -> barrier
  combine    the down GroupGEMM's epilogue stores each segment into the SOURCE rank's return buffer (MxbProblem.C =
             peer pointer, c_row = its first row there)                                        -> barrier
  then the local top-k weighted sum.  Capacity is the worst case (W * T * k rows), so routing can never overflow a peer.

exchange="nccl" (the baseline the push path is measured against, and the path of gloo/CPU host-logic tests):
`all_to_all_single` of the quantized rows + scales and of the fp16 down rows with per-peer split sizes; its one host
synchronisation per step is the read-back of the [W, E] count matrix the split sizes come from.

Overlap: the replicated shared expert depends on the local tokens only, so its chain (act-quant, gate_up GEMM,
Hadamard+quant, down GEMM) runs on a side stream through a second handle whose persistent kernels leave `comm_sms`
SMs free (mxb_set_sm_limit) for the dispatch kernel (push) or NCCL's kernels: a persistent GEMM CTA owns its SM's
register file, hence the SM split rather than plain stream concurrency.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import ctypes
import os

import numpy as np
import torch
import torch.distributed as dist

from . import ops
from .config import QLayerConfig
from .layer import MoELayer


def expert_bounds(E: int, W: int) -> List[int]:
    """Contiguous expert slices: rank r owns [b[r], b[r+1])."""
    return [r * E // W for r in range(W + 1)]


def balanced_expert_order(freq: Sequence[float], W: int) -> List[int]:
    """Expert placement from an access-frequency profile (the `access_freq` of the reference's gate trace,
    moe_tracer.py:189-206): longest-processing-time assignment into the W contiguous slices of expert_bounds — most
    popular expert first, each to the least loaded rank that still has a free slot.  Returns the expert order; rank r owns
    order[bounds[r]:bounds[r+1]].  With contiguous placement the step waits for the rank that holds the popular experts."""
    E = len(freq)
    bounds = expert_bounds(E, W)
    cap = [bounds[r + 1] - bounds[r] for r in range(W)]
    load = [0.0] * W
    slots: List[List[int]] = [[] for _ in range(W)]
    for e in sorted(range(E), key=lambda i: (-float(freq[i]), i)):
        r = min((r for r in range(W) if len(slots[r]) < cap[r]), key=lambda r: (load[r], r))
        slots[r].append(e)
        load[r] += float(freq[e])
    return [e for r in range(W) for e in sorted(slots[r])]


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
                 group: Optional[dist.ProcessGroup] = None, weights: Optional[dict] = None, overlap_shared: bool = True,
                 comm_sms: int = 20, exchange: str = "auto", expert_order: Optional[Sequence[int]] = None,
                 share: Optional["EPMoELayer"] = None):
        """expert_order: placement of the experts (default 0..E-1): rank r owns expert_order[bounds[r]:bounds[r+1]]
        (balanced_expert_order builds one from a popularity profile).  Internally experts are relabelled by their position
        in the order, so every kernel still sees contiguous slices; topk_ids keep the caller's numbering.
        share: an earlier layer of the same stack (same shapes, same world): activation, exchange (symmetric-memory) and routing
        buffers, the side stream and the side handle are re-used — the layers of a stack run one after the other."""
        self.group = group
        self.W = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.bounds = expert_bounds(E, self.W)
        order = list(expert_order) if expert_order is not None else list(range(E))
        assert sorted(order) == list(range(E)), "expert_order must be a permutation of the experts"
        self.order = order
        mine = order[self.bounds[self.rank]:self.bounds[self.rank + 1]]
        self.T_local = max_tokens
        if exchange == "auto":
            exchange = "push" if (self.W > 1 and dist.get_backend(group) == "nccl") else "nccl"
        self.exchange = exchange
        self.push = exchange == "push"
        # receive capacity in rows: `capacity` x the balanced share (NCCL path, checked on the host every step);
        # the push path cannot check before it writes into a peer, so it holds the worst case
        self.cap_rows = max_tokens * topk * self.W if self.push else int(max_tokens * topk * capacity) + 128
        self.q_all = [(qlayer.experts[str(e)].gate.a_bits, qlayer.experts[str(e)].gate.a_gsize) for e in order]  # by position
        self.sender_quant = all(a < 16 for a, _ in self.q_all)
        self.any_g128 = any(g == 128 for a, g in self.q_all if a < 16)
        cap_tokens = (self.cap_rows + topk - 1) // topk
        # MoELayer seeds its generator per expert, so the model is the same whatever W is (and the replicated shared expert
        # has the same weights on every rank)
        self._scale_cols = H // 128 if self.any_g128 else 1
        super().__init__(handle, E, topk, H, I, ns, qlayer, max_tokens=cap_tokens, hadamard_down=hadamard_down,
                         expert_ids=mine, seed=seed, weight_scale=weight_scale, weights=weights, share=share)
        d = self.dev
        relabel = np.empty(E, dtype=np.int32)
        relabel[np.asarray(order)] = np.arange(E, dtype=np.int32)
        self.relabel = None if order == list(range(E)) else torch.from_numpy(relabel).to(d)
        R = max_tokens * topk
        cols = self._scale_cols
        B = self._buf
        if not self.push and self.sender_quant and self.x_sc.shape[1] != cols:  # receive buffer must match the senders' scale width
            self.x_sc = B("x_sc_recv", (self.x_sc.shape[0], cols), torch.float16)
        nccl = not self.push  # staging buffers of the NCCL exchange
        self.send_q8 = B("send_q8", (R, H), torch.int8) if (nccl and self.sender_quant) else None
        self.send_sc = B("send_sc", (R, cols), torch.float16) if (nccl and self.sender_quant) else None
        self.send_f16 = None if (self.sender_quant or not nccl) else B("send_f16", (R, H), torch.float16)
        self.recv_f16 = None if (self.sender_quant or not nccl) else B("recv_f16", (self.cap_rows, H), torch.float16)
        self.seg_offs = B("seg_offs", (self.W * self.EL + 1,), torch.int32, zero=True)
        if self.push:
            self._setup_push(R, H, cols)
        else:
            self.back = B("back", (R, H), torch.float16)
            self.counts_all = B("counts_all", (self.W, E), torch.int32, zero=True)
        self._last_seg = None
        self._build_segment_problems()
        # the shared expert's chain on a side stream, on a handle whose persistent kernels leave SMs to the exchange
        if os.environ.get("MXB_EP_OVERLAP") is not None:
            overlap_shared = os.environ["MXB_EP_OVERLAP"] != "0"
        comm_sms = int(os.environ.get("MXB_EP_COMM_SMS", comm_sms))
        self.tail_wait = os.environ.get("MXB_EP_TAIL_WAIT", "0") != "0"
        self.overlap = bool(overlap_shared and ns and self.W > 1)
        if self.overlap and share is not None and getattr(share, "overlap", False):
            self.h_side, self.side = share.h_side, share.side
            self.ev_in, self.ev_mid, self.ev_gu, self.ev_side = share.ev_in, share.ev_mid, share.ev_gu, share.ev_side
        elif self.overlap:
            self.h_side = ops.Handle(handle.device)
            self.h_side.set_sm_limit(handle.sm_count - comm_sms)
            self.side = torch.cuda.Stream(d)
            self.ev_in, self.ev_mid, self.ev_gu, self.ev_side = (torch.cuda.Event() for _ in range(4))
        # the dispatch kernel is NVLink-bound: when the shared expert runs beside it, it keeps to the SMs left free
        self.push_ctas = int(os.environ.get("MXB_EP_PUSH_CTAS", comm_sms * 8 if self.overlap else 0))
        # MXB_EP_TIMING=1: CUDA events at the phase boundaries of forward(), read with timing_report()
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

    def _symm(self, name: str, shape, dtype):
        """A symmetric-memory buffer + its peer views (collective: every rank calls it in the same order); pooled like
        MoELayer._buf so that the layers of a stack rendezvous once."""
        import torch.distributed._symmetric_memory as symm
        key = ("symm", name, tuple(int(v) for v in shape), dtype)
        ent = self._pool.get(key)
        if ent is None:
            gname = (self.group or dist.group.WORLD).group_name
            t = symm.empty(key[2], dtype=dtype, device=self.dev)
            t.zero_()
            hdl = symm.rendezvous(t, gname)
            peers = [t if p == self.rank else hdl.get_buffer(p, key[2], dtype) for p in range(self.W)]
            ent = (t, peers, hdl)
            self._pool[key] = ent
        return ent

    def _alloc(self, T: int):
        """push exchange: the receive buffers of the gate_up GEMM live in symmetric memory (the peers' dispatch kernels
        write them); they are created first and handed to the parent's allocation under their names."""
        self.peer_q8 = self.peer_f16 = self.peer_sc = None
        if self.push:
            cap, H, cols = self.cap_rows, self.H, self._scale_cols
            ov = {"x_q8": None, "x_sc": None, "x_f16": None}
            # decided from ALL experts' formats, not the local ones: the rendezvous is collective and senders quantize per expert
            if any(a < 16 for a, _ in self.q_all):
                ov["x_q8"], self.peer_q8, _ = self._symm("x_q8", (cap, H), torch.int8)
                ov["x_sc"], self.peer_sc, _ = self._symm("x_sc", (cap, cols), torch.float16)
            if any(a >= 16 for a, _ in self.q_all):
                ov["x_f16"], self.peer_f16, _ = self._symm("x_f16", (cap, H), torch.float16)
            self._override = ov
        super()._alloc(T)
        if self.push:  # local experts may use fewer formats than the senders: keep every receive buffer reachable
            self.x_q8, self.x_sc, self.x_f16 = self._override["x_q8"], self._override["x_sc"], self._override["x_f16"]

    def _setup_push(self, R: int, H: int, cols: int):
        """Return buffer, count matrix and the device-side plan of the push exchange."""
        W, E = self.W, self.E
        fresh = ("symm", "counts_all", (W, E), torch.int32) not in self._pool
        self.back, self.peer_back, _ = self._symm("back", (R, H), torch.float16)
        self.counts_all, self.peer_counts, self.sync = self._symm("counts_all", (W, E), torch.int32)
        self.back_row = self._buf("back_row", (W * self.EL,), torch.int32, zero=True)
        self.dest = self._buf("dest", (E * ctypes.sizeof(ops.MxbRowDest),), torch.uint8, zero=True)
        self.scale_cols = cols
        if fresh:
            torch.cuda.synchronize()
            dist.barrier(group=self.group)

    def _build_segment_problems(self):
        H, I = self.H, self.I
        R = self.x_q8.shape[0] if self.x_q8 is not None else self.x_f16.shape[0]
        gu, dn = [], []
        for s in range(self.W):
            for i, ew in enumerate(self.experts[:self.EL]):
                mr = self.seg_offs[s * self.EL + i: s * self.EL + i + 2]
                gu.append(self._problem(ew.q_gu, self.x_q8, self.x_f16, self.x_sc, ew.gu_packed, ew.gu_scale, self.gu_out, R, 2 * I, H, mr,
                                        flags=ops.EPI_SILU_MUL if self.fused_silu else 0))
                if self.push:  # results go straight into the source rank's return buffer, at the segment's row there
                    j = s * self.EL + i
                    dn.append(self._problem(ew.q_dn, self.h_q8, self.h_f16, self.h_sc, ew.dn_packed, ew.dn_scale, self.peer_back[s],
                                            R, H, I, mr, c_row=self.back_row[j:j + 1]))
                else:
                    dn.append(self._problem(ew.q_dn, self.h_q8, self.h_f16, self.h_sc, ew.dn_packed, ew.dn_scale, self.dn_out, R, H, I, mr))
        self._gu_routed, self._dn_routed = gu, dn
        self._seg_actq_gu = [(e.q_gu.a_bits, e.q_gu.a_gsize) for _ in range(self.W) for e in self.experts[:self.EL]]
        self._seg_actq_dn = [(e.q_dn.a_bits, e.q_dn.a_gsize) for _ in range(self.W) for e in self.experts[:self.EL]]

    def _finalize_problems(self, T: int, x: torch.Tensor):
        if not getattr(self, "overlap", False):
            return super()._finalize_problems(T, x)
        s = self.experts[-1]
        key = (T, x.data_ptr() if s.q_gu.a_bits >= 16 else 0)
        if self._T_built == key:
            return
        if not hasattr(self, "_lists"):
            self._lists = {}
        if key not in self._lists:
            Is = self.I * self.ns
            A_f16 = x if s.q_gu.a_bits >= 16 else None
            sgu = [self._problem(s.q_gu, self.sx_q8, A_f16, self.sx_sc, s.gu_packed, s.gu_scale, self.s_gu, T, 2 * Is, self.H, None,
                                 flags=ops.EPI_SILU_MUL if self.fused_silu else 0)]
            sdn = [self._problem(s.q_dn, self.sh_q8, self.sh_f16, self.sh_sc, s.dn_packed, s.dn_scale, self.s_dn, T, self.H, Is, None)]
            gu, dn = list(self._gu_routed), list(self._dn_routed)
            if len(self._lists) >= 8:
                self._lists.clear()
            self._lists[key] = (self.h.pack_problems(gu), self.h.pack_problems(dn), self.h.pack_problems(sgu), self.h.pack_problems(sdn),
                                (gu, dn, sgu, sdn))
        self._gu_arr, self._dn_arr, self._sgu_arr, self._sdn_arr, self._keep = self._lists[key]
        self._T_built = key

    def _shared_gate_up(self, x: torch.Tensor, T: int):
        """side stream: act-quant + gate_up GEMM of the shared expert over the local tokens"""
        hs, s = self.h_side, self.experts[-1]
        if s.q_gu.a_bits < 16:
            hs.quant_act(x, None, T, None, [(s.q_gu.a_bits, s.q_gu.a_gsize)], False, self.sx_q8, None, self.sx_sc)
        hs.group_gemm(self._sgu_arr)

    def _shared_down(self, T: int):
        hs, s = self.h_side, self.experts[-1]
        aq = [(s.q_dn.a_bits, s.q_dn.a_gsize)]
        if self.fused_silu:
            if not self.sh_direct:
                hs.quant_act(self.s_gu, None, T, None, aq, self.hadamard_down, self.sh_q8, self.sh_f16, self.sh_sc)
        else:
            hs.silu_mul_quant(self.s_gu, T, None, aq, self.hadamard_down, self.sh_q8, self.sh_f16, self.sh_sc)
        hs.group_gemm(self._sdn_arr)

    def _a2a(self, out: torch.Tensor, inp: torch.Tensor, out_splits, in_splits):
        dist.all_to_all_single(out, inp, output_split_sizes=out_splits, input_split_sizes=in_splits, group=self.group)

    def forward(self, x: torch.Tensor, topk_ids: torch.Tensor, topk_w: torch.Tensor, out: Optional[torch.Tensor] = None):
        if self.relabel is not None:  # caller's expert ids -> positions in the placement order
            topk_ids = self.relabel[topk_ids.long()]
        if self.push:
            return self._forward_push(x, topk_ids, topk_w, out)
        return self._forward_nccl(x, topk_ids, topk_w, out)

    def _forward_push(self, x: torch.Tensor, topk_ids: torch.Tensor, topk_w: torch.Tensor, out: Optional[torch.Tensor] = None):
        h, T, k, E, W = self.h, x.shape[0], self.topk, self.E, self.W
        n_send = T * k
        cur = torch.cuda.current_stream()
        self._finalize_problems(T, x)
        self._mark("start")
        offs, src, slot = h.permute(topk_ids, E)
        h.ep_push_counts(offs, E, self.peer_counts, self.rank)
        self.sync.barrier(channel=0)
        h.ep_plan(self.counts_all, W, self.rank, E, self.peer_q8, self.peer_f16, self.peer_sc, self.seg_offs, self.back_row, self.dest)
        if self.overlap:  # shared gate_up next to the dispatch
            self.ev_in.record(cur)
            self.side.wait_event(self.ev_in)
            with torch.cuda.stream(self.side):
                self._shared_gate_up(x, T)
                if not self.tail_wait:
                    self._shared_down(T)
                    self.ev_side.record(self.side)
        self._mark("permute+counts+plan")
        h.quant_act_push(x, src, n_send, offs, self.q_all, self.dest, self.scale_cols, self.push_ctas)
        self.sync.barrier(channel=1)
        self._mark("dispatch_push")
        self._run_segments(x, self.cap_rows, T)  # row counts live on the device (seg_offs): kernels stop at the last offset
        self._mark("experts+combine_push")
        if self.overlap and self.tail_wait:  # shared down next to the tail of the exchange
            self.ev_mid.record(cur)
            self.side.wait_event(self.ev_mid)
            with torch.cuda.stream(self.side):
                self._shared_down(T)
                self.ev_side.record(self.side)
        self.sync.barrier(channel=2)
        if self.overlap:
            cur.wait_event(self.ev_side)
        self._mark("barrier+shared_tail")
        y = h.combine(self.back, slot, topk_w, self.s_dn if self.ns else None, out)
        self._mark("combine")
        return y

    def _forward_nccl(self, x: torch.Tensor, topk_ids: torch.Tensor, topk_w: torch.Tensor, out: Optional[torch.Tensor] = None):
        h, T, k, E, W = self.h, x.shape[0], self.topk, self.E, self.W
        n_send = T * k
        cur = torch.cuda.current_stream()
        self._finalize_problems(T, x)
        self._mark("start")
        offs, src, slot = h.permute(topk_ids, E)
        my_counts = (offs[1:] - offs[:-1]).contiguous()
        # sender-side quantization needs the local offsets only: issue it before the count exchange
        if self.sender_quant:
            h.quant_act(x, src, n_send, offs, self.q_all, False, self.send_q8, None, self.send_sc)
        else:
            h.quant_act(x, src, n_send, None, [(16, -1)], False, None, self.send_f16, None)  # gather rows into sorted order
        dist.all_gather_into_tensor(self.counts_all.view(-1), my_counts, group=self.group)
        if self.overlap:  # shared gate_up under the count read-back and the dispatch
            self.ev_in.record(cur)
            self.side.wait_event(self.ev_in)
            with torch.cuda.stream(self.side):
                self._shared_gate_up(x, T)
                self.ev_gu.record(self.side)
        self._mark("permute+quant+allgather")
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
            self._a2a(self.x_q8[:n_recv], self.send_q8[:n_send], recv, send)
            self._a2a(self.x_sc[:n_recv], self.send_sc[:n_send], recv, send)
        else:
            self._a2a(self.recv_f16[:n_recv], self.send_f16[:n_send], recv, send)
            h.quant_act(self.recv_f16, None, n_recv, self.seg_offs, self._seg_actq_gu, False, self.x_q8, self.x_f16, self.x_sc)
        self._mark("dispatch_a2a")
        # ---- local experts on (source, expert) segments (+ the replicated shared expert when it is not overlapped)
        self._run_segments(x, n_recv, T)
        self._mark("experts")
        if self.overlap:  # shared down under the combine exchange
            self.ev_mid.record(cur)
            self.side.wait_event(self.ev_mid)
            with torch.cuda.stream(self.side):
                self._shared_down(T)
                self.ev_side.record(self.side)
        # ---- combine
        self._a2a(self.back[:n_send], self.dn_out[:n_recv], send, recv)
        self._mark("combine_a2a")
        if self.overlap:
            cur.wait_event(self.ev_side)
        y = h.combine(self.back, slot, topk_w, self.s_dn if self.ns else None, out)
        self._mark("combine")
        return y

    def _run_segments(self, x: torch.Tensor, n_recv: int, T: int):
        h = self.h
        self._finalize_problems(T, x)
        shared_here = self.ns and not getattr(self, "overlap", False)
        if shared_here:
            s = self.experts[-1]
            if s.q_gu.a_bits < 16:
                h.quant_act(x, None, T, None, [(s.q_gu.a_bits, s.q_gu.a_gsize)], False, self.sx_q8, None, self.sx_sc)
        h.group_gemm(self._gu_arr)
        if self.fused_silu:  # the gate_up epilogue already wrote silu(gate)*up
            if not self.h_direct:
                h.quant_act(self.gu_out, None, n_recv, self.seg_offs, self._seg_actq_dn, self.hadamard_down, self.h_q8, self.h_f16,
                            self.h_sc)
            if shared_here and not self.sh_direct:
                s = self.experts[-1]
                h.quant_act(self.s_gu, None, T, None, [(s.q_dn.a_bits, s.q_dn.a_gsize)], self.hadamard_down, self.sh_q8, self.sh_f16,
                            self.sh_sc)
        else:
            h.silu_mul_quant(self.gu_out, n_recv, self.seg_offs, self._seg_actq_dn, self.hadamard_down, self.h_q8, self.h_f16, self.h_sc)
            if shared_here:
                s = self.experts[-1]
                h.silu_mul_quant(self.s_gu, T, None, [(s.q_dn.a_bits, s.q_dn.a_gsize)], self.hadamard_down, self.sh_q8, self.sh_f16,
                                 self.sh_sc)
        h.group_gemm(self._dn_arr)

    # ------------------------------------------------------------------ bench helpers
    def gemm_only(self):
        """Re-issue every GroupGEMM launch of the last step on its buffers: the routed segment lists AND, when the
        replicated shared expert runs through its own lists (overlap), those too — the launches gemm_ops() counts."""
        self.h.group_gemm(self._gu_arr)
        self.h.group_gemm(self._dn_arr)
        if getattr(self, "overlap", False):
            self.h.group_gemm(self._sgu_arr)
            self.h.group_gemm(self._sdn_arr)

    def last_counts(self):
        """rows per local expert processed in the last step (summed over sources)"""
        if self.push:
            o = self.seg_offs.cpu().numpy().astype(np.int64)
            return (o[1:] - o[:-1]).reshape(self.W, self.EL).sum(0).tolist()
        if self._last_seg is None:
            return [0] * self.EL
        return self._last_seg.reshape(self.W, self.EL).sum(0).tolist()

    def launches_per_step(self) -> int:
        perm1 = 1 if self.T_local * self.topk <= 2048 else 0  # permute_small_kernel: one launch instead of two
        if self.push:  # permute (2), counts, plan, quant+push, GroupGEMM x2, (rotate+)quant, combine; shared: quant, GEMM x2, quant
            n = 9 - perm1 + (4 if self.ns and self.overlap else (2 if self.ns else 0))
            if self.fused_silu:
                n -= (1 if self.h_direct else 0) + (1 if self.ns and self.sh_direct else 0)
            return n
        n = 8 - perm1 + (2 if self.ns else 0) + (0 if self.sender_quant else 1)
        if self.fused_silu:
            n -= (1 if self.h_direct else 0) + (1 if self.ns and self.sh_direct else 0)
        return n
