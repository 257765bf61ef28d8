#!/usr/bin/env python
"""Kernel-only sweep of the persistent GroupGEMM: one dense problem per kind / tile N, CUDA-event timed.
Usage (GPU box): python tools/gemm_sweep.py [M N K]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mxmoe_b200 import ops  # noqa: E402


def main():
    M, N, K = (int(v) for v in sys.argv[1:4]) if len(sys.argv) >= 4 else (8192, 11264, 2048)
    h = ops.Handle(0)
    dev = "cuda"
    C = torch.empty(M, N, dtype=torch.float16, device=dev)
    sa = torch.ones(M, dtype=torch.float16, device=dev)
    sb = torch.ones(N, dtype=torch.float16, device=dev)
    A8 = torch.randint(-7, 8, (M, K), device=dev, dtype=torch.int8)
    B8 = torch.randint(-127, 128, (N, K), device=dev, dtype=torch.int8)
    B4 = torch.randint(0, 256, (N, K // 2), device=dev, dtype=torch.uint8)
    A16 = (torch.rand(M, K, device=dev) - 0.5).half()
    B16 = (torch.rand(N, K, device=dev) - 0.5).half()
    sbw = torch.ones(K // 128 * N * 2, dtype=torch.float16, device=dev)
    cases = []
    # tile_n = 0 on an eligible problem -> CTA-pair kernel (cta_group::2)
    cases.append(("w8a8 pair", ops.Problem(A8, B8, C, M, N, K, 8, 8, -1, True, scale_a=sa, scale_b=sb)))
    cases.append(("w4a4 pair", ops.Problem(A8, B4, C, M, N, K, 4, 4, -1, True, scale_a=sa, scale_b=sb)))
    cases.append(("fp16 pair", ops.Problem(A16, B16, C, M, N, K)))
    for bn in (256, 128, 64):
        cases.append((f"w8a8 bn{bn}", ops.Problem(A8, B8, C, M, N, K, 8, 8, -1, True, scale_a=sa, scale_b=sb, tile_n=bn)))
    for bn in (256, 128, 64):
        cases.append((f"w4a4 bn{bn}", ops.Problem(A8, B4, C, M, N, K, 4, 4, -1, True, scale_a=sa, scale_b=sb, tile_n=bn)))
    for bn in (256, 128):
        cases.append((f"fp16 bn{bn}", ops.Problem(A16, B16, C, M, N, K, tile_n=bn)))
    cases.append(("w4a16_g128 bn128", ops.Problem(A16, B4, C, M, N, K, 4, 16, 128, False, scale_b=sbw, tile_n=128)))
    sag = torch.ones(M, K // 128, dtype=torch.float16, device=dev)
    sbg = torch.ones(K // 128, N, dtype=torch.float16, device=dev)
    cases.append(("w4a4_g128 bn64", ops.Problem(A8, B4, C, M, N, K, 4, 4, 128, True, scale_a=sag, scale_b=sbg)))
    ops_total = 2.0 * M * N * K
    for name, p in cases:
        arr = h.pack_problems([p])
        for _ in range(3):
            h.group_gemm(arr)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 10
        e0.record()
        for _ in range(n):
            h.group_gemm(arr)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        print(f"{name:18s} M={M} N={N} K={K}  {ms*1e3:9.1f} us  {ops_total/ms/1e9:8.1f} TOPS", flush=True)
    # library yardstick (not the product): torch fp16 matmul of the same shape
    for _ in range(3):
        torch.matmul(A16, B16.T)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        torch.matmul(A16, B16.T)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"{'cuBLAS fp16 (yardstick)':18s} {ms*1e3:9.1f} us  {ops_total/ms/1e9:8.1f} TFLOPS")


if __name__ == "__main__":
    main()
