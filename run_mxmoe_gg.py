#!/usr/bin/env python
"""Drop-in for the reference driver `run_mxmoe_gg.py` (ref run_mxmoe_gg.py:32-141): same command line, same
workload JSON (out/workloads/...) and the same per-GroupGEMM CSV (`kernel_name,avg_time,TFLOPS,speedup`,
ref mxmoe/kernels/src/test.cu:858-864) — but no source generation and no cmake: the problems of the workload
are handed to the precompiled sm_100a library (libmxmoe_b200.so) as run-time descriptors.

Synthetic data exactly like the reference bench (ref test.cu:877-915): A and W ~ U(-1,1) fp16, seed 42;
quantized operands are produced by the library's own quantize/pack kernels (the reference feeds raw fp16 bytes,
test.cu:518-519).  `speedup` is relative to a torch fp16 matmul loop over the same problems on the same GPU
(the reference's column is relative to CUTLASS sm80 grouped GEMM).
"""
import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from mxmoe_b200 import config as C  # noqa: E402
from mxmoe_b200 import workload as W  # noqa: E402


def qstr_of(p) -> str:
    if p["w_bits"] == 16 and p["a_bits"] == 16:
        return "fp16"
    return f"w{p['w_bits']}a{p['a_bits']}_g{p['gsize']}_{'sym' if p['sym'] else 'asym'}"


def build_problems(h, ops, torch, plist, seed=42, tile_n_of=None):
    """One Problem per workload entry with M > 0 (empty problems are skipped like ref_bind.cu:360).
    tile_n_of: optional qstr -> MxbProblem.tile_n (from --tile_config; 0 = the library's own choice)."""
    g = torch.Generator(device="cuda").manual_seed(seed)
    probs, keep, total_ops = [], [], 0
    for p in plist:
        M, N, K = p["shape"]
        if M == 0:
            continue
        w_bits, a_bits, gsize, sym = p["w_bits"], p["a_bits"], p["gsize"], bool(p["sym"])
        A = (torch.rand(M, K, generator=g, device="cuda") * 2 - 1).half()
        Wt = (torch.rand(N, K, generator=g, device="cuda") * 2 - 1).half()
        Cb = torch.empty(M, N, dtype=torch.float16, device="cuda")
        tn = tile_n_of(qstr_of(p)) if tile_n_of else 0
        if w_bits == 16 and a_bits == 16:
            probs.append(ops.Problem(A, Wt, Cb, M, N, K, tile_n=tn))
            keep += [A, Wt, Cb]
        else:
            packed, sc = h.quant_pack_weight(Wt, w_bits, a_bits, gsize, sym)
            if a_bits < 16:
                q8, _, sa = h.quant_act(A, None, M, None, [(a_bits, gsize)])
                probs.append(ops.Problem(q8, packed, Cb, M, N, K, w_bits, a_bits, gsize, True, scale_a=sa, scale_b=sc, tile_n=tn))
                keep += [q8, sa]
            else:
                probs.append(ops.Problem(A, packed, Cb, M, N, K, w_bits, 16, gsize, sym, scale_b=sc, tile_n=tn))
                keep += [A]
            keep += [packed, sc, Cb]
            del Wt
        total_ops += 2 * M * N * K
    return probs, keep, total_ops


def time_ms(torch, fn, warmup=10, iters=30):
    """median of CUDA-event timings (ref test_utils.h:97-191: warm-ups, then the median of the timed runs)"""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return statistics.median(ts)


def main():
    ap = argparse.ArgumentParser(description="Bench workloads for the groupgemm kernel (B200-native library).")
    ap.add_argument("--model", type=str, default="qwen2_moe", help="Model ID.")
    ap.add_argument("--dataset", type=str, default="wiki2", help="Dataset ID.")
    ap.add_argument("--qconfig", type=str, default=None, help="Path to the quantization config file.")
    ap.add_argument("--tile_config", type=str, default=None, help="Path to the tile config file.")
    ap.add_argument("--qstr", type=str, default=None, help="Short string to represent quantization config.")
    ap.add_argument("--bs", type=int, default=512, help="Batch size.")
    ap.add_argument("--layer", type=int, default=-1, help="Layer index.")
    # additions: the reference reads calib/gate/.../moe-gate.json, which is not shipped
    ap.add_argument("--trace", type=str, default=None, help="moe-gate.json gate trace (moe_tracer.py:174-211 schema).")
    ap.add_argument("--routing", type=str, default="fixture", choices=["fixture", "uniform", "zipf"],
                    help="synthetic trace when --trace is not given")
    ap.add_argument("--dry-run", action="store_true", help="only write the workload JSON (no GPU needed)")
    args = ap.parse_args()

    if args.qconfig is not None:
        qcfg_list = C.get_qcfg_list(args.qconfig, args.layer)
        suffix = "-" + args.qconfig.split("wbits")[1].split(".json")[0] + ".json" if "wbits" in args.qconfig else "-qcfg.json"
        qconfig = json.load(open(args.qconfig))
    elif args.qstr is not None:
        qcfg_list, suffix, qconfig = [args.qstr], f"-{args.qstr}.json", None
    else:
        qcfg_list, suffix, qconfig = ["fp16"], "-fp16.json", None
    layers = list(range(C.MODEL_ID_TO_LAYERS[args.model])) if args.layer == -1 else [args.layer]
    if args.model not in C.MODEL_SHAPES:
        raise SystemExit(f"unknown model {args.model}")
    tiles = C.load_tile_config(args.tile_config) if args.tile_config else {}
    print(f"qconfigs: {sorted(qcfg_list)}" + (f"; tile_config entries: {len(tiles)}" if tiles else ""))

    workload_path = os.path.join(ROOT, "out", "workloads", f"{args.model}-{args.dataset}-{args.bs}{suffix}")
    if args.trace:
        trace = json.load(open(args.trace))
    else:
        kind = args.routing if (args.routing != "fixture" or args.model == "qwen2_moe") else "uniform"
        trace = W.synthetic_gate_trace(args.model, layers, kind)
    wl = {"num_tokens": args.bs}
    for L in layers:
        one = W.generate_workload_from_gate_trace(trace, args.bs, L, qconfig=qconfig, qstr=args.qstr)
        wl.update({k: v for k, v in one.items() if k.startswith("layer-")})
    W.save_workload(wl, workload_path)
    print(f"Save generated workloads to `{workload_path}`")
    if args.dry_run:
        return

    import torch
    from mxmoe_b200 import ops
    h = ops.Handle(torch.cuda.current_device())
    os.makedirs(os.path.join(ROOT, "out", "bench"), exist_ok=True)
    for L in layers:
        print(f"Processing Layer {L}...")
        # --tile_config: the entry of every qconfig decides the kernel path / UMMA N of its problems (config.tile_hint)
        hints = {q: C.tile_hint(C.tile_for(tiles, q, L, qcfg_list)) for q in qcfg_list}
        if tiles:
            print("  tile_config -> tile_n per qconfig (0 = selector's choice):", hints)
        for part in ("gate_up", "down"):
            probs, keep, total_ops = build_problems(h, ops, torch, wl[f"layer-{L}"][part], tile_n_of=lambda q: hints.get(q, 0))
            choice = h.select_tiles(probs)
            paths = sorted({(c.path, c.tile_m, c.tile_n) for c in choice})
            print(f"  {part}: kernel path / tile per problem kind: {paths}")
            arr = h.pack_problems(probs)
            ms = time_ms(torch, lambda: h.group_gemm(arr))
            As = [(torch.rand(p.M, p.K, device="cuda") - 0.5).half() for p in probs]
            Bs = [(torch.rand(p.N, p.K, device="cuda") - 0.5).half() for p in probs]
            ms_ref = time_ms(torch, lambda: [torch.matmul(a, b.T) for a, b in zip(As, Bs)], warmup=3, iters=10)
            out = os.path.join(ROOT, "out", "bench", f"{args.model}-{args.dataset}-{args.bs}{suffix[:-5]}-layer-{L}-{part}.csv")
            with open(out, "w") as f:
                f.write("kernel_name,avg_time,TFLOPS,speedup\n")
                f.write(f"torch_fp16_matmul_loop,{ms_ref:.6f},{total_ops / ms_ref / 1e9:.3f},1.000\n")
                f.write(f"mxb_group_gemm_sm100a[{'+'.join(sorted(qcfg_list))}],{ms:.6f},{total_ops / ms / 1e9:.3f},{ms_ref / ms:.3f}\n")
            print(f"  {part}: {len(probs)} problems, {ms * 1e3:.1f} us, {total_ops / ms / 1e9:.1f} TOPS -> {out}")
            del probs, keep, arr, As, Bs
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
