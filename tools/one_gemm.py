#!/usr/bin/env python
"""One dense GroupGEMM problem, a few launches — the target of `ncu` captures.
Usage: python tools/one_gemm.py KIND [M N K] [reps]   KIND = w8a8 | w4a4 | fp16, kernel chosen by MXB_GEMM_PATH."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mxmoe_b200 import ops  # noqa: E402

kind = sys.argv[1]
M, N, K = (int(v) for v in sys.argv[2:5]) if len(sys.argv) >= 5 else (8192, 11264, 2048)
reps = int(sys.argv[5]) if len(sys.argv) >= 6 else 5
h = ops.Handle(0)
C = torch.empty(M, N, dtype=torch.float16, device="cuda")
sa = torch.ones(M, dtype=torch.float16, device="cuda")
sb = torch.ones(N, dtype=torch.float16, device="cuda")
if kind == "w8a8":
    p = ops.Problem(torch.randint(-127, 128, (M, K), device="cuda", dtype=torch.int8),
                    torch.randint(-127, 128, (N, K), device="cuda", dtype=torch.int8), C, M, N, K, 8, 8, -1, True, scale_a=sa, scale_b=sb)
elif kind == "w4a4":
    p = ops.Problem(torch.randint(-7, 8, (M, K), device="cuda", dtype=torch.int8),
                    torch.randint(0, 256, (N, K // 2), device="cuda", dtype=torch.uint8), C, M, N, K, 4, 4, -1, True, scale_a=sa, scale_b=sb)
else:
    p = ops.Problem((torch.rand(M, K, device="cuda") - 0.5).half(), (torch.rand(N, K, device="cuda") - 0.5).half(), C, M, N, K)
arr = h.pack_problems([p])
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
h.group_gemm(arr)
torch.cuda.synchronize()
e0.record()
for _ in range(reps):
    h.group_gemm(arr)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print(f"{kind} M={M} N={N} K={K}: {ms*1e3:.1f} us, {2.0*M*N*K/ms/1e9:.1f} TOPS")
