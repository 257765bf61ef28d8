#!/usr/bin/env python
"""Same-box GPU baseline: the REFERENCE's own GroupGEMM kernel (oracle/_ref/libref_gg.so = its cta_gemm / mm_tile templates
compiled for sm_100a, oracle/ref_gg_harness.cu) against mxb_group_gemm on the problem lists of a bench workload.

    python tools/ref_kernel_compare.py [--workload c2] [--iters 20]  > profiles/ref_kernel_compare_rNN.jsonl

Both sides run PLAIN GroupGEMMs over the same shapes and bit-widths ([M, N] fp16 outputs; no fused silu epilogue on our
side), operands in each side's own packed layout, contents random (timing does not depend on values; parity on real
quantized operands is tests/test_gpu_ref_kernel.py).  Median of CUDA-event timings like the reference bench
(test_utils.h:128-189).  This is measurement tooling, not part of bench.py (which may not execute oracle/ on the GPU)."""
import argparse
import ctypes as C
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


class Dim3(C.Structure):
    _fields_ = [("x", C.c_uint32), ("y", C.c_uint32), ("z", C.c_uint32)]


class QParams(C.Structure):
    _fields_ = [("a_bits", C.c_int32), ("w_bits", C.c_int32), ("gsize", C.c_int32), ("sym", C.c_uint8), ("pad", C.c_uint8 * 3)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    import torch
    import bench
    from mxmoe_b200 import config as Cfg
    from mxmoe_b200 import ops
    lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_gg.so"))
    lib.ref_gg_launch.restype = C.c_int
    lib.ref_gg_launch.argtypes = [C.c_void_p] * 9 + [C.c_int, C.c_void_p, C.c_int, C.c_void_p]
    wl = bench.WORKLOADS[a.workload]
    s = Cfg.MODEL_SHAPES[wl["model"]]
    ql = bench.qlayer_for(wl["model"], wl["layer"], wl["q"])
    T = wl["bs"]
    ids, _ = bench.routing(wl["model"], T, seed=0, device=None)
    counts = torch.bincount(ids.flatten().long(), minlength=s["E"]).tolist()
    h = ops.Handle(0)
    g = torch.Generator(device="cuda").manual_seed(0)

    def rnd8(*shape):
        return torch.randint(-128, 128, shape, generator=g, device="cuda", dtype=torch.int16).to(torch.int8)

    for part in ("gate_up", "down"):
        units = [(e, counts[e], s["I"]) for e in range(s["E"]) if counts[e] > 0] + ([(s["E"], T, s["I"] * s["ns"])] if s["ns"] else [])
        ours, keep, sizes, qps, rA, rB, rsa, rsb, rC = [], [], [], [], [], [], [], [], []
        tot = 0
        for (e, M, I) in units:
            q = ql.experts[str(e)].gate if (part == "gate_up" or e == s["E"]) else ql.experts[str(e)].down
            N, K = (2 * I, s["H"]) if part == "gate_up" else (s["H"], I)
            bits = q.w_bits
            assert bits in (4, 8, 16) and q.a_bits == bits
            tot += 2 * M * N * K
            Cb = torch.empty(M, N, dtype=torch.float16, device="cuda")
            Cr = torch.empty(M, N, dtype=torch.float16, device="cuda")
            if bits == 16:
                A = (torch.rand(M, K, generator=g, device="cuda") - 0.5).half()
                B = (torch.rand(N, K, generator=g, device="cuda") - 0.5).half()
                ours.append(ops.Problem(A, B, Cb, M, N, K))
                ra, rb, sa, sb = A, B, None, None
            else:
                A8 = rnd8(M, K) if bits == 8 else torch.randint(-7, 8, (M, K), generator=g, device="cuda", dtype=torch.int16).to(torch.int8)
                Bn = rnd8(N, K * bits // 8)
                sa = (torch.rand(M, generator=g, device="cuda") * 0.01 + 0.001).half()
                sb = (torch.rand(N, generator=g, device="cuda") * 0.01 + 0.001).half()
                ours.append(ops.Problem(A8, Bn, Cb, M, N, K, bits, bits, -1, True, scale_a=sa, scale_b=sb))
                ra, rb = rnd8(M, K * bits // 8), rnd8(N, K * bits // 8)  # reference layout: K*bits/16 packed halves per row
            keep += [Cb, Cr, ra, rb, sa, sb]
            rA.append(ra); rB.append(rb); rsa.append(sa); rsb.append(sb); rC.append(Cr)
            sizes.append(Dim3(M, N, K)); qps.append(QParams(bits, bits, -1, 1))
        P = len(units)
        hs, hq = (Dim3 * P)(*sizes), (QParams * P)(*qps)
        d_sizes = torch.frombuffer(bytearray(bytes(hs)), dtype=torch.uint8).cuda()
        d_q = torch.frombuffer(bytearray(bytes(hq)), dtype=torch.uint8).cuda()
        d_prefix = torch.zeros(P, dtype=torch.int32, device="cuda")
        tab = lambda ts: torch.tensor([t.data_ptr() if t is not None else 0 for t in ts], dtype=torch.int64, device="cuda")
        tabs = [tab(t) for t in (rA, rB, rsa, rsb, rC)]
        arr = h.pack_problems(ours)

        def run_ref():
            rc = lib.ref_gg_launch(*[t.data_ptr() for t in tabs], d_sizes.data_ptr(), C.addressof(hs), d_q.data_ptr(), C.addressof(hq), P,
                                   d_prefix.data_ptr(), 0, None)
            assert rc == 0, rc

        def timed(fn):
            for _ in range(5):
                fn()
            torch.cuda.synchronize()
            ts = []
            for _ in range(a.iters):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); fn(); e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            return statistics.median(ts)

        with torch.cuda.stream(torch.cuda.default_stream()):  # the reference launcher is hard-wired to stream 0 semantics
            ms_ref = timed(run_ref)
        ms_ours = timed(lambda: h.group_gemm(arr))
        print(json.dumps({"workload": a.workload, "list": part, "problems": P, "ops": tot,
                          "reference_kernel_ms": ms_ref, "reference_kernel_tops": tot / ms_ref / 1e9,
                          "mxb_group_gemm_ms": ms_ours, "mxb_group_gemm_tops": tot / ms_ours / 1e9, "speedup": ms_ref / ms_ours,
                          "reference_kernel": "cta_gemm_multistage_v2 / _qab_v2, tiles 128x128x{64,128,128}, mma.sync (s4 emulated on sm_100), "
                                              "grid = #SM, 4 warps", "note": "plain GroupGEMM on both sides, random operand bytes"}))
        del ours, keep, rA, rB, rsa, rsb, rC, arr
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
