#!/usr/bin/env python
"""One shape of quant_act+Hadamard, a few launches (ncu target)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mxmoe_b200 import ops
h = ops.Handle(0)
rows, I = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (32768, 1408)
x = (torch.rand(rows, I, device="cuda") * 4 - 2).half()
q8 = torch.empty(rows, I, dtype=torch.int8, device="cuda"); sc = torch.empty(rows, 1, dtype=torch.float16, device="cuda")
for _ in range(4):
    h.quant_act(x, None, rows, None, [(8, -1)], True, q8, None, sc)
torch.cuda.synchronize()
