#!/usr/bin/env python
"""B200 cost table for the bit-width solver (replaces the RTX-4090 perf/performance_table.json of the reference, which
get_runtime_cost reads at bits_solver.py:518-542): for every strategy and a ladder of expert sizes, the MEASURED time of
the expert's gate_up and down GroupGEMM problems next to the selector's estimate (mxb_select_tiles), which is what
mxmoe_b200.bits_solver uses.  Each problem is timed in a list of 8 identical experts (one launch), CUDA events.
Usage (GPU box): python tools/cost_table.py [model] > profiles/cost_table_b200_r01.json"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mxmoe_b200 import config as C  # noqa: E402
from mxmoe_b200 import ops  # noqa: E402

STRATEGIES = ["fp16", "w8a8_g-1_sym", "w8a8_g-1_sym_E4M3", "w4a4_g-1_sym", "w4a4_g128_sym", "w4a16_g128_asym", "w4a16_g-1_sym"]
ROWS = [8, 64, 136, 546, 2048, 8192]
COPIES = 8


def time_list(h, arr, n=10):
    for _ in range(3):
        h.group_gemm(arr)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        h.group_gemm(arr)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3  # us


def main():
    model = sys.argv[1] if len(sys.argv) > 1 else "qwen2_moe"
    s = C.MODEL_SHAPES[model]
    H, I = s["H"], s["I"]
    h = ops.Handle(0)
    dev = "cuda"
    table = {"model": model, "gpu": torch.cuda.get_device_name(0), "copies_per_launch": COPIES, "unit": "us per expert problem",
             "strategies": {}}
    for qs in STRATEGIES:
        q = C.QLinearConfig.from_qstr(qs)
        ent = {}
        for blk, (N, K) in (("gate_up", (2 * I, H)), ("down", (H, I))):
            Wt = ((torch.rand(N, K, device=dev) * 2 - 1) * 0.05).half()
            if q.w_bits == 16:
                B, sb = Wt, None
            else:
                B, sb = h.quant_pack_weight(Wt, q.w_bits, q.a_bits, q.w_gsize, q.w_sym)
            for M in ROWS:
                A16 = (torch.rand(M, K, device=dev) - 0.5).half()
                if q.a_bits >= 16:
                    A, sa = A16, None
                else:
                    A, _, sa = h.quant_act(A16, None, M, None, [(q.a_bits, q.a_gsize)])
                probs = []
                for _ in range(COPIES):
                    Cb = torch.empty(M, N, dtype=torch.float16, device=dev)
                    probs.append(ops.Problem(A, B, Cb, M, N, K, q.w_bits, q.a_bits, q.w_gsize, q.w_sym if q.w_bits < 16 else True,
                                             scale_a=sa, scale_b=sb))
                us = time_list(h, h.pack_problems(probs)) / COPIES
                ch = h.select_tiles(probs[:1])[0]
                ent[f"{blk}:{M}"] = {"shape": [M, N, K], "measured_us": round(us, 2), "selector_est_us": round(float(ch.est_us), 2),
                                     "path": int(ch.path), "tile": [int(ch.tile_m), int(ch.tile_n), int(ch.tile_k)],
                                     "tops": round(2.0 * M * N * K / us / 1e6, 1)}
        table["strategies"][qs] = ent
    print(json.dumps(table, indent=1))


if __name__ == "__main__":
    main()
