"""The reference's CPU torch path for the MoE-layer GroupGEMMs, timed as the CPU baseline.  TEST/BENCH
INFRASTRUCTURE ONLY (bench.py's cpu_baseline / --impl reference legs).

What the reference executes per expert on the accuracy side: weights fake-quantized offline
(Quantizer.fake_quant, mxmoe/quant/quant.py:104-106, :460-461), every Linear input fake-quantized per token
on the fly (mlp_inp_quant_hook, quant.py:126-130), then torch.nn.functional.linear.  Restated here with the
oracle's quant_minmax port for the numerics and torch fp32 matmuls on all host threads (fp16 matmul on CPU
is unrepresentatively slow)."""
import os
import time

import numpy as np
import torch

from . import oracle as O


def _fq_act(x: np.ndarray, a_bits: int, gsize: int) -> torch.Tensor:
    if a_bits >= 16:
        return torch.from_numpy(x.astype(np.float32))
    return torch.from_numpy(O.fake_quant(x, a_bits, gsize, True).astype(np.float32))


def prepare_expert(Wgu, Wdn, q_gu, q_dn):
    """offline: fake-quantized fp32 weights.  q = (w_bits, a_bits, gsize, sym)"""
    def fq(W, q):
        return torch.from_numpy((W if q[0] >= 16 else O.fake_quant(W, q[0], q[2], q[3])).astype(np.float32))
    return fq(Wgu, q_gu), fq(Wdn, q_dn)


def expert_forward(x: np.ndarray, Wgu_f: torch.Tensor, Wdn_f: torch.Tensor, q_gu, q_dn) -> torch.Tensor:
    a = _fq_act(x, q_gu[1], q_gu[2])
    gu = torch.nn.functional.linear(a, Wgu_f).half()
    I = gu.shape[1] // 2
    h = (torch.nn.functional.silu(gu[:, :I].float()).half().float() * gu[:, I:].float()).half()
    a2 = _fq_act(h.numpy(), q_dn[1], q_dn[2])
    return torch.nn.functional.linear(a2, Wdn_f).half()


def time_sample(problems, threads=None, repeats=3):
    """problems: list of (M, H, I, q_gu, q_dn).  Returns (ops, best seconds, threads)."""
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    rng = np.random.default_rng(42)
    prepared = []
    ops = 0
    for (M, H, I, q_gu, q_dn) in problems:
        Wgu = ((rng.random((2 * I, H)) * 2 - 1) * 0.05).astype(np.float16)
        Wdn = ((rng.random((H, I)) * 2 - 1) * 0.05).astype(np.float16)
        x = (rng.random((M, H)) * 2 - 1).astype(np.float16)
        prepared.append((x, *prepare_expert(Wgu, Wdn, q_gu, q_dn), q_gu, q_dn))
        ops += 2 * M * (2 * I) * H + 2 * M * H * I
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        for (x, a, b, q1, q2) in prepared:
            expert_forward(x, a, b, q1, q2)
        best = min(best, time.perf_counter() - t0)
    return ops, best, threads
