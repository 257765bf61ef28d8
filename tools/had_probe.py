#!/usr/bin/env python
"""Times the (Hadamard +) per-token quantization kernels alone on the c2 shapes, for the fused-silu pipeline
(input = silu(gate)*up fp16 [rows, I]) and the two-step one (input = gate|up [rows, 2I]).
MXB_HAD_DBG masks phases of the mma.sync kernel (profiling experiment)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mxmoe_b200 import ops
h = ops.Handle(0)
def t(f):
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 10 * 1e3
for rows, I in ((32768, 1408), (8192, 5632)):
    gu = (torch.rand(rows, 2 * I, device="cuda") * 4 - 2).half()
    x = (torch.rand(rows, I, device="cuda") * 4 - 2).half()
    q8 = torch.empty(rows, I, dtype=torch.int8, device="cuda"); sc = torch.empty(rows, 1, dtype=torch.float16, device="cuda")
    for had in (True, False):
        a = t(lambda: h.silu_mul_quant(gu, rows, None, [(8, -1)], had, q8, None, sc))
        b = t(lambda: h.quant_act(x, None, rows, None, [(8, -1)], had, q8, None, sc))
        mb = rows * I * 3 / 1e6
        print(f"dbg={os.environ.get('MXB_HAD_DBG','0')} rows={rows} I={I} hadamard={had}: silu_mul_quant {a:.1f} us, quant_act {b:.1f} us "
              f"({mb / b * 1e3 / 1e3:.0f} GB/s of {mb:.0f} MB)")
