"""The reference's CPU torch path for the MoE-layer GroupGEMMs, timed as the CPU baseline.  TEST/BENCH
INFRASTRUCTURE ONLY (bench.py's cpu_baseline / --impl reference legs).

What the reference executes per expert on the accuracy side: weights fake-quantized offline
(Quantizer.fake_quant, mxmoe/quant/quant.py:104-106, :460-461), every Linear input fake-quantized per token
on the fly (mlp_inp_quant_hook, quant.py:126-130), the down projection's input rotated by the online Hadamard
hook (rotation.py:403-421 -> matmul_hadU, hadamard_utils.py:62-83; the CUDA FHT extension cannot run on a CPU,
the file's own pure-torch matmul_hadU can), then torch.nn.functional.linear.

Everything inside the timed region is torch on all host threads, in fp32 (fp16 matmul on CPU is
unrepresentatively slow): `fake_quant_t` is a line-for-line torch port of quant_minmax + dequant
(quant.py:40-106: abs().amax / div / sub / clamp / round / mul / add), `hadamard_t` of matmul_hadU.  The numpy
oracle (oracle.py) is only used outside the timed region: offline weight fake-quant and the table of hadK signs.
"""
import os
import statistics
import time

import numpy as np
import torch

from . import oracle as O


def fake_quant_t(t: torch.Tensor, bits: int, gsize: int, sym: bool) -> torch.Tensor:
    """Quantizer.fake_quant (quant.py:40-106) on a torch tensor, clip = (1, 1)."""
    if bits >= 16:
        return t
    if bits == O.BITS_E4M3:  # the FP8 kind (oracle.quant_e4m3): per-row scale absmax / 448, codes on torch's own float8_e4m3fn grid
        scale = (t.abs().amax(dim=-1, keepdim=True) / O.E4M3_MAX).to(torch.float16).to(t.dtype)
        scale = torch.where(scale == 0, torch.ones_like(scale), scale)
        return (t / scale).clamp(-O.E4M3_MAX, O.E4M3_MAX).to(torch.float8_e4m3fn).to(t.dtype) * scale
    g = t.shape[-1] if gsize == -1 else gsize
    r = t.reshape(-1, g)
    if sym:
        upper = (1 << (bits - 1)) - 1
        lower = -upper
        scale = r.abs().amax(dim=-1, keepdim=True) / upper
        zp = 0
    else:
        upper, lower = (1 << bits) - 1, 0
        gmin, gmax = r.aminmax(dim=-1, keepdim=True)
        scale = (gmax - gmin) / upper
        zp = gmin
    q = r.sub(zp).div(scale).clamp(lower, upper).round()
    return q.mul(scale).add(zp).reshape(t.shape)


_HADK = {}


def hadamard_t(X: torch.Tensor) -> torch.Tensor:
    """Online Hadamard of the down projection's input: matmul_hadU_cuda semantics (hadamard_utils.py:96-105) with the
    butterfly loop of matmul_hadU (:62-83) for the power-of-two part, fp32."""
    n = X.shape[-1]
    if n not in _HADK:
        hk, K = O.get_hadK(n)
        _HADK[n] = (None if hk is None else torch.from_numpy(np.asarray(hk, dtype=np.float32)), K)
    hk, K = _HADK[n]
    m = n // K
    y = X.reshape(-1, K, m)
    h = 1
    while h < m:  # Sylvester butterflies over the last axis
        y = y.reshape(-1, K, m // (2 * h), 2, h)
        y = torch.stack((y[:, :, :, 0, :] + y[:, :, :, 1, :], y[:, :, :, 0, :] - y[:, :, :, 1, :]), dim=3)
        h *= 2
    y = y.reshape(-1, K, m)
    if K > 1:
        y = hk.view(1, K, K) @ y
    return (y / float(np.sqrt(n))).reshape(X.shape)


def prepare_expert(Wgu, Wdn, q_gu, q_dn):
    """offline: fake-quantized fp32 weights.  q = (w_bits, a_bits, gsize, sym)"""
    def fq(W, q):
        return torch.from_numpy((W if q[0] >= 16 else O.fake_quant(W, q[0], q[2], q[3])).astype(np.float32))
    return fq(Wgu, q_gu), fq(Wdn, q_dn)


def expert_forward(x: torch.Tensor, Wgu_f: torch.Tensor, Wdn_f: torch.Tensor, q_gu, q_dn, hadamard: bool = False) -> torch.Tensor:
    """x fp32 [M, H] -> fp32 [M, H]; activation hooks as quant.py:126-130, rotation hook as rotation.py:403-421."""
    a = fake_quant_t(x, q_gu[1], q_gu[2], True)
    gu = torch.nn.functional.linear(a, Wgu_f)
    I = gu.shape[1] // 2
    h = torch.nn.functional.silu(gu[:, :I]) * gu[:, I:]
    if hadamard:
        h = hadamard_t(h)
    a2 = fake_quant_t(h, q_dn[1], q_dn[2], True)
    return torch.nn.functional.linear(a2, Wdn_f)


class CpuSample:
    """A prepared sample (offline weight fake-quant done once) whose `run()` is one timed pass."""

    def __init__(self, problems, threads=None, hadamard=False, matmul_only=False):
        self.threads = threads or os.cpu_count()
        torch.set_num_threads(self.threads)
        rng = np.random.default_rng(42)
        self.prepared, self.ops = [], 0
        self.hadamard = hadamard and not matmul_only
        for (M, H, I, q_gu, q_dn) in problems:
            Wgu = ((rng.random((2 * I, H)) * 2 - 1) * 0.05).astype(np.float16)
            Wdn = ((rng.random((H, I)) * 2 - 1) * 0.05).astype(np.float16)
            x = torch.from_numpy((rng.random((M, H)) * 2 - 1).astype(np.float32))
            if matmul_only:
                q_gu, q_dn = (q_gu[0], 16, q_gu[2], q_gu[3]), (q_dn[0], 16, q_dn[2], q_dn[3])
            self.prepared.append((x, *prepare_expert(Wgu, Wdn, q_gu, q_dn), q_gu, q_dn))
            self.ops += 2 * M * (2 * I) * H + 2 * M * H * I

    def run(self) -> float:
        """one pass over the sample; seconds"""
        with torch.no_grad():
            t0 = time.perf_counter()
            for (x, a, b, q1, q2) in self.prepared:
                expert_forward(x, a, b, q1, q2, self.hadamard)
            return time.perf_counter() - t0


def time_sample(problems, threads=None, repeats=3, hadamard=False, matmul_only=False):
    """problems: list of (M, H, I, q_gu, q_dn).  Returns (ops, median seconds, threads).
    matmul_only: the same problems without the activation hooks (fake-quant, Hadamard) — the reported side figure."""
    smp = CpuSample(problems, threads, hadamard, matmul_only)
    smp.run()  # first touch of the weights / thread pool start-up
    return smp.ops, statistics.median(smp.run() for _ in range(repeats)), smp.threads
