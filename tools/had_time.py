#!/usr/bin/env python
"""Time quant_act+Hadamard alone (CUDA events, 20 launches after 5) for a list of rows x n shapes; GB/s = (2 n + n + 2) bytes per row."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mxmoe_b200 import ops
h = ops.Handle(0)
shapes = [(32768, 1408), (8192, 5632), (65536, 2560), (8192, 20480), (16384, 14336), (8192, 7168), (8192, 10240)]
if len(sys.argv) > 2:
    shapes = [(int(sys.argv[1]), int(sys.argv[2]))]
for rows, n in shapes:
    x = (torch.rand(rows, n, device="cuda") * 4 - 2).half()
    q8 = torch.empty(rows, n, dtype=torch.int8, device="cuda"); sc = torch.empty(rows, 1, dtype=torch.float16, device="cuda")
    for _ in range(5):
        h.quant_act(x, None, rows, None, [(8, -1)], True, q8, None, sc)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(20):
        h.quant_act(x, None, rows, None, [(8, -1)], True, q8, None, sc)
    b.record(); torch.cuda.synchronize()
    us = a.elapsed_time(b) * 1000 / 20
    print(f"hadamard+quant {rows} x {n}: {us:8.1f} us  {rows * (3 * n + 2) / us * 1e-3:7.0f} GB/s  env PIPE_WIDE={os.environ.get('MXB_HAD_PIPE_WIDE', '1')}", flush=True)
    del x, q8, sc
