#!/usr/bin/env python
"""bench.py — MoE-layer GroupGEMM throughput (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c1|c3|c4|c5|c5_stack]

A step = one pass of the whole hot path over one batch of synthetic tokens: permute -> act-quant ->
GroupGEMM(gate_up) -> silu*mul(+Hadamard)+quant -> GroupGEMM(down) -> combine.  `value` = sum 2*M*N*K of the
step's GroupGEMM problems (the reference's TFLOPS numerator, test.cu:109-116) / step time, inputs resident in
HBM; `e2e` = the same through host buffers (pinned H2D of tokens+routing, D2H of the layer output) each step.
N > 1: expert-parallel, weak scaling (every rank brings its own `bs` tokens; dispatch/combine pushed over NVLink by the
quant kernel / the down GEMM epilogue).  After the timed region every rank checks a sample of its output rows against
the single-GPU layer (`parity_check`); a mismatch aborts with a non-zero exit code.  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # BASELINE.json configs[1]: the configuration the metric is quoted on
    "c2": dict(model="qwen2_moe", layer=11, bs=8192, q="mixed5", hadamard=True,
               name="Qwen1.5-MoE-A2.7B layer-11 mixed w4a4_g-1_sym+w8a8_g-1_sym (wbits 5.0, +Hadamard) bs=8192"),
    "c1": dict(model="qwen2_moe", layer=11, bs=512, q="fp16", hadamard=False,
               name="Qwen1.5-MoE-A2.7B layer-11 FP16 bs=512"),
    "c3": dict(model="ds2", layer=1, bs=64, q="w4a16_g128_asym", hadamard=False,
               name="DeepSeek-V2-Lite MoE layer w4a16_g128 weight-only bs=64"),
    "c4": dict(model="mixtral", layer=0, bs=8192, q="w8a8_g-1_sym", hadamard=False,
               name="Mixtral-8x7B MoE layer w8a8_g-1_sym bs=8192"),
    # the reference's FP8 kind (tile_config.py:45) on the c4 shape: tcgen05.mma kind::f8f6f4
    "c4_e4m3": dict(model="mixtral", layer=0, bs=8192, q="w8a8_g-1_sym_E4M3", hadamard=False,
                    name="Mixtral-8x7B MoE layer w8a8_g-1_sym_E4M3 bs=8192"),
    "c5": dict(model="qwen2_moe_57b", layer=0, bs=8192, q="mixed5", hadamard=True,
               name="Qwen2-57B-A14B MoE layer mixed w4a4+w8a8 bs=8192"),
    # BASELINE.json configs[4]: every MoE block of the model, each with its own solver-assigned qconfig
    "c5_stack": dict(model="qwen2_moe_57b", layer=0, layers=list(range(28)), bs=8192, q="ilp5", hadamard=True,
                     name="Qwen2-57B-A14B full MoE stack: 28 blocks, per-block ILP qconfig (mxmoe_b200.bits_solver, w4a4+w8a8, "
                          "wbits 5.0, r 0.75, B200 cost table), +Hadamard, bs=8192"),
}
ILP_QCONFIG = {"qwen2_moe_57b": os.path.join(ROOT, "mxmoe_b200", "data", "qconfig_qwen2_moe_57b_wbits5.0_r0.75.json")}
_ILP_CACHE = {}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm_gbs=d["hbm_gbs"], bf16=d["bf16_tflops"], bf16_sustained=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    src="measured")
    return dict(hbm_gbs=6650.0, bf16=1590.0, bf16_sustained=1400.0, src="fallback")


def qlayer_for(model, layer, q):
    from mxmoe_b200 import config as C
    from mxmoe_b200 import workload as W
    s = C.MODEL_SHAPES[model]
    if q == "mixed5":
        return C.qmodel_cfg_from_dict(W.synthetic_mixed_qconfig(model, [layer], 5.0, seed=0)).layers[str(layer)]
    if q == "ilp5":
        # the qconfig the repo's own solver wrote for this model (python -m mxmoe_b200.bits_solver --model qwen2_moe_57b --wbits 5.0
        # --batch 8192 --r 0.75 --layers 0..27 --cost_table profiles/cost_table_b200_r01.json), the reference's JSON schema
        if model not in _ILP_CACHE:
            _ILP_CACHE[model] = C.qmodel_cfg_from_dict({k: v for k, v in json.load(open(ILP_QCONFIG[model])).items() if k != "LT"})
        return _ILP_CACHE[model].layers[str(layer)]
    units = s["E"] + (1 if s["ns"] else 0)
    return C.build_uni_qlayer_cfg(C.QLinearConfig.from_qstr(q), units)


def routing(model, T, seed, device):
    """Synthetic top-k routing: k distinct experts per token, drawn from the expert-popularity profile of the
    reference's shipped workload fixture (qwen2_moe) or uniform; softmax-like positive weights."""
    import numpy as np
    import torch
    from mxmoe_b200 import config as C
    from mxmoe_b200 import workload as W
    s = C.MODEL_SHAPES[model]
    rng = np.random.default_rng(seed)
    kind = "fixture" if model == "qwen2_moe" else "uniform"
    freq = np.array(W.synthetic_gate_trace(model, [0], kind)["layer-0"]["access_freq"], dtype=np.float64)
    p = freq / freq.sum()
    g = -np.log(-np.log(rng.random((T, s["E"])))) + np.log(p)[None, :]  # Gumbel top-k == sampling without replacement
    ids = np.argsort(-g, axis=1)[:, :s["topk"]].astype(np.int32)
    w = rng.random((T, s["topk"])).astype(np.float32) + 0.1
    w /= w.sum(1, keepdims=True)
    return torch.from_numpy(ids), torch.from_numpy(w)


class ClockSampler(threading.Thread):
    """SM clock, power and throttle reasons during a timed region (B200_PROFILING.md recipe) through NVML: one read right
    before the region (`start()` returns after it), reads every 2 ms while it runs, one read right after (`finish()`), so that
    even a region of a few hundred microseconds carries a clock record."""
    NAMES = None

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.power, self.reasons, self.max_mhz = index, False, [], [], set(), None
        self.nv = self.h = None
        try:
            import pynvml as nv
            nv.nvmlInit()
            self.nv, self.h = nv, nv.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
            self.names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                          nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                          nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                          nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        except Exception as e:  # NVML missing: report it, do not fake
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def read(self):
        if self.nv is None:
            return
        try:
            self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
            self.power.append(self.nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
            r = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            for bit, name in self.names.items():
                if r & bit:
                    self.reasons.add(name)
        except Exception as e:
            self.reasons.add(f"nvml_error:{type(e).__name__}")

    def begin(self):
        self.read()
        self.start()
        return self

    def run(self):
        while not self.stop_flag:
            self.read()
            time.sleep(0.002)

    def finish(self):
        self.stop_flag = True
        self.join(timeout=2)
        self.read()
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_min_mhz": (s[0] if s else None), "sm_max_mhz": self.max_mhz,
                "power_w_max": (round(max(self.power), 1) if self.power else None), "samples": len(s), "reasons": sorted(self.reasons)}


def cpu_sample_problems(model, layer, q, n=3):
    """A bounded sample of the workload for the CPU arm: the first n routed experts at M = bs*k/E."""
    from mxmoe_b200 import config as C
    s = C.MODEL_SHAPES[model]
    ql = qlayer_for(model, layer, q)
    t = lambda c: (c.w_bits, c.a_bits, c.w_gsize, c.w_sym if c.w_bits < 16 else True)
    out = []
    for e in range(n):
        qe = ql.experts[str(e)]
        out.append((e, t(qe.gate), t(qe.down)))
    shared = None
    if s["ns"]:
        qe = ql.experts[str(s["E"])]
        shared = (t(qe.gate), t(qe.gate))  # the shared down problem takes the gate config (gen_workload.py:101)
    return s, out, shared


def cpu_sample(wl, n_experts):
    """The reference's CPU torch path (oracle/cpu_path.py) on a bounded sample that keeps the workload's mix: n routed
    experts at the mean row count, the shared expert on the matching share of tokens (rows_routed / top-k), the online
    Hadamard when the workload has it."""
    s, sample, shared = cpu_sample_problems(wl["model"], wl["layer"], wl["q"], n_experts)
    M = max(1, wl["bs"] * s["topk"] // s["E"])
    probs = [(M, s["H"], s["I"], qg, qd) for (_, qg, qd) in sample]
    Ms = 0
    if shared is not None:
        Ms = max(1, n_experts * M // s["topk"])
        probs.append((Ms, s["H"], s["I"] * s["ns"], shared[0], shared[1]))
    what = (f"{n_experts} routed experts of the workload (M={M} each)"
            + (f" + the shared expert on {Ms} tokens (the routed:shared ratio of the full step)" if Ms else "")
            + ": per-token fake-quant hook (torch port of quant.py:40-106)"
            + (" + online Hadamard hook (torch matmul_hadU)" if wl["hadamard"] else "")
            + " + F.linear, gate_up -> silu*mul -> down, torch fp32 on all host threads")
    return probs, what


def run_cpu(wl, n_experts, repeats):
    """cpu_baseline of the default arm: median of `repeats` passes + the same sample without the activation hooks."""
    from oracle import cpu_path
    probs, what = cpu_sample(wl, n_experts)
    ops, sec, threads = cpu_path.time_sample(probs, repeats=repeats, hadamard=wl["hadamard"])
    base = dict(value=ops / sec / 1e12, unit="TOPS", cores=threads, kind="port", sample=what + f", median of {repeats}")
    ops2, sec2, _ = cpu_path.time_sample(probs, repeats=repeats, matmul_only=True)
    base["matmul_only_value"] = ops2 / sec2 / 1e12  # what the host cores do on the matmuls alone
    return base


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--bs", type=int, default=None)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    wl = dict(WORKLOADS[args.workload])
    if args.bs:
        wl["bs"] = args.bs
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    W = max(1, args.gpus)
    steps, warmup = args.steps, max(3, args.warmup)
    cfg = {"workload": wl["name"], "bs_per_gpu": wl["bs"], "routing": "synthetic top-k (fixture expert popularity), seed 0",
           "weights": ("U(-1,1)*0.05 fp16, seed 42, fake-quantized offline (host)" if args.impl == "reference" else
                       "U(-1,1)*0.05 fp16, seed 42, quantized+packed on device"), "parallelism": f"ep{W}" if W > 1 else "single",
           "l2": "working set per step (packed weights + activations, >0.9 GB) exceeds the 126 MB L2; no explicit flush"}

    if args.impl == "reference":
        if rank != 0:
            return
        # the reference's CPU torch path on this box's host cores: every step is one pass over a bounded sample
        from oracle import cpu_path
        probs, what = cpu_sample(wl, 3)
        smp = cpu_path.CpuSample(probs, hadamard=wl["hadamard"])
        for _ in range(warmup):
            smp.run()
        times = []
        for _ in range(steps):
            times.append(smp.run())
            if sum(times) > 150:  # keep the arm within a few minutes on a slow host
                break
        steps = len(times)
        t_tot = sum(times)
        v = smp.ops * steps / t_tot / 1e12
        base = dict(value=v, unit="TOPS", cores=smp.threads, kind="port", sample=what + f", {steps} timed passes")
        print(json.dumps({"impl": "reference", "metric": "MoE-layer GroupGEMM TOPS", "value": v, "unit": "TOPS", "n_gpus": W,
                          "steps": steps, "warmup": warmup, "ms_per_step": 1e3 * t_tot / steps, "higher_is_better": True,
                          "scaling": "weak", "vs_baseline": None, "dtype": "f32 (host) over int8/int4 grids", "data": "synthetic",
                          "config": cfg, "cpu_baseline": base,
                          "e2e": {"value": v, "unit": "TOPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    import torch
    import torch.distributed as dist
    from mxmoe_b200 import config as C
    from mxmoe_b200 import ops
    from mxmoe_b200.layer import HostPipeline, MoELayer, MoEStack

    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    s = C.MODEL_SHAPES[wl["model"]]
    T = wl["bs"]
    layer_ids = wl.get("layers", [wl["layer"]])
    nL = len(layer_ids)
    qls = [qlayer_for(wl["model"], li, wl["q"]) for li in layer_ids]
    h = ops.Handle(local)
    order = None
    if world > 1:
        from mxmoe_b200.ep import EPMoELayer, balanced_expert_order
        from mxmoe_b200 import workload as Wl
        # expert placement from the same popularity profile the routing is drawn from (the reference's gate-trace access_freq)
        kind = "fixture" if wl["model"] == "qwen2_moe" else "uniform"
        freq = Wl.synthetic_gate_trace(wl["model"], [0], kind)["layer-0"]["access_freq"]
        order = balanced_expert_order(freq, world) if os.environ.get("MXB_EP_PLACEMENT", "contiguous") == "balanced" else None
        cfg["placement"] = "popularity-balanced (LPT over the gate trace's access_freq)" if order is not None else "contiguous"
    blocks = []
    for j, (li, ql) in enumerate(zip(layer_ids, qls)):
        share = blocks[0] if blocks else None
        if world > 1:
            blocks.append(EPMoELayer(h, s["E"], s["topk"], s["H"], s["I"], s["ns"], ql, max_tokens=T, hadamard_down=wl["hadamard"],
                                     weight_scale=0.05, expert_order=order, seed=42 + li, share=share))
        else:
            blocks.append(MoELayer(h, s["E"], s["topk"], s["H"], s["I"], s["ns"], ql, max_tokens=T, hadamard_down=wl["hadamard"],
                                   weight_scale=0.05, seed=42 + li, share=share))
    stacked = nL > 1
    layer = MoEStack(blocks) if stacked else blocks[0]
    g = torch.Generator(device="cpu").manual_seed(1234 + rank)
    x_host = (torch.rand(T, s["H"], generator=g) * 2 - 1).half().pin_memory()
    routes = [routing(wl["model"], T, seed=rank + 1000 * j, device=dev) for j in range(nL)]
    if stacked:
        ids_host, w_host = torch.stack([r[0] for r in routes]).contiguous(), torch.stack([r[1] for r in routes]).contiguous()
    else:
        ids_host, w_host = routes[0]
    ids_host, w_host = ids_host.pin_memory(), w_host.pin_memory()
    x, ids, w = x_host.to(dev), ids_host.to(dev), w_host.to(dev)
    out = torch.empty(T, s["H"], dtype=torch.float16, device=dev)
    out_host = torch.empty(T, s["H"], dtype=torch.float16).pin_memory()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # decode-sized steps are launch-bound: the step is captured once as a CUDA graph (MoELayer.capture) and replayed
    use_graph = world == 1 and T <= 256 and not stacked
    if use_graph:
        graph = layer.capture(x, ids, w, out)
        step = graph.replay
        cfg["step"] = "CUDA graph replay of layer.forward (decode)"
    else:
        step = lambda: layer.forward(x, ids, w, out)
    for _ in range(warmup):
        step()
    if world > 1 and os.environ.get("MXB_EP_TIMING", "0") == "1":
        for b in blocks:
            b.timing_report()  # drop the warm-up marks
    sampler = ClockSampler(local).begin()
    ms = timed(step, steps)
    clocks = sampler.finish()

    if world > 1 and os.environ.get("MXB_EP_TIMING", "0") == "1":
        rep = blocks[0].timing_report()
        allrep = [None] * world
        dist.all_gather_object(allrep, {k: round(v, 4) for k, v in rep.items()})
        if rank == 0:  # per-rank phase timelines of the first block (mean ms per phase)
            for r_, d_ in enumerate(allrep):
                print(f"ep phases rank {r_} (ms):", d_, file=sys.stderr)
    # ops actually executed (from this step's routing), summed over blocks and ranks
    block_ops = []
    for j, b in enumerate(blocks):
        if stacked:  # the row counts of block j: re-run its permute bookkeeping from the routing (offs is shared by the stack)
            b.forward(x, ids[j], w[j], out)
        block_ops.append(float(b.gemm_ops(b.last_counts(), T)))
    my_gemm_ops = sum(block_ops)
    my_ops = torch.tensor([my_gemm_ops], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(my_ops)
    total_ops = float(my_ops.item())
    value = total_ops * steps / (ms * 1e-3) / 1e12

    per_block = None
    if stacked:  # one more pass with an event after every block
        marks = []
        barrier()
        e0 = torch.cuda.Event(enable_timing=True)
        e0.record()
        layer.forward(x, ids, w, out, marks=marks)
        torch.cuda.synchronize()
        prev, per_block = e0, []
        for ev in marks:
            per_block.append(round(prev.elapsed_time(ev), 4))
            prev = ev

    # ---- parity of what was just timed (N > 1): a sample of this rank's output rows against the single-GPU layer
    parity = None
    if world > 1:
        n_chk = min(T, 256)
        li, ql = layer_ids[-1], qls[-1]
        ref = MoELayer(ops.Handle(local), s["E"], s["topk"], s["H"], s["I"], s["ns"], ql, max_tokens=n_chk, hadamard_down=wl["hadamard"],
                       weight_scale=0.05, seed=42 + li)
        ids_l, w_l = (ids[-1], w[-1]) if stacked else (ids, w)
        y_ep = blocks[-1].forward(x, ids_l, w_l, out)[:n_chk].clone()
        y_ref = ref.forward(x[:n_chk].contiguous(), ids_l[:n_chk].contiguous(), w_l[:n_chk].contiguous())
        torch.cuda.synchronize()
        same = bool(torch.equal(y_ep.view(torch.int16), y_ref.view(torch.int16))) and bool(torch.isfinite(y_ep.float()).all()) \
            and bool(y_ep.float().abs().max() > 0)
        flag = torch.tensor([1 if same else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        parity = {"rows": n_chk * world, "against": "single-GPU MoELayer with the same per-expert weights, same tokens and routing "
                  "(itself checked against the oracle at this size: tests/test_gpu_layer_configs.py)", "bit_identical": bool(flag.item())}
        del ref
        torch.cuda.empty_cache()

    # ---- sustained: the same step looped for >= 3 s (the headline above is a burst of `steps` steps at boost clocks)
    sustained = None
    if os.environ.get("MXB_BENCH_SUSTAINED", "1") != "0":
        n_sus = max(steps, int(3000.0 / max(ms / steps, 1e-3)) + 1)
        smp2 = ClockSampler(local).begin()
        ms_sus = timed(step, n_sus)
        ck2 = smp2.finish()
        v_sus = total_ops * n_sus / (ms_sus * 1e-3) / 1e12
        sustained = {"seconds": ms_sus * 1e-3, "steps": n_sus, "ms_per_step": ms_sus / n_sus, "value": v_sus, "unit": "TOPS", "clocks": ck2}

    # e2e: pinned host buffers in, pinned host buffer out, every step, through the public host-batch API
    # (mxmoe_b200.layer.HostPipeline: upload / kernels / download on three streams, two device slots)
    pipe = HostPipeline(layer, T, graph=use_graph)
    out_hosts = [out_host, torch.empty_like(out_host).pin_memory()]
    n_sub = [0]
    def step_e2e():
        pipe.submit(x_host, ids_host, w_host, out_hosts[n_sub[0] & 1])
        n_sub[0] += 1
    for _ in range(3):
        step_e2e()
    pipe.join()
    def run_e2e(n):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            step_e2e()
        pipe.join()  # the timed region ends when the last step's result is in host memory
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    ms_e2e = run_e2e(steps)
    e2e = {"value": total_ops * steps / (ms_e2e * 1e-3) / 1e12, "unit": "TOPS",
           "h2d_bytes_per_step": x_host.numel() * 2 + ids_host.numel() * 4 + w_host.numel() * 4,
           "d2h_bytes_per_step": out_host.numel() * 2, "ms_per_step": ms_e2e / steps}
    if world > 1 and os.environ.get("MXB_BENCH_COPY_ONLY", "1") != "0":
        # what the host side alone can do at this world size: the same pinned H2D + D2H per step, no kernels in between
        s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        def copies():
            with torch.cuda.stream(s_in):
                x.copy_(x_host, non_blocking=True); ids.copy_(ids_host, non_blocking=True); w.copy_(w_host, non_blocking=True)
            with torch.cuda.stream(s_out):
                out_host.copy_(out, non_blocking=True)
        def copies_join():
            torch.cuda.current_stream().wait_stream(s_in); torch.cuda.current_stream().wait_stream(s_out)
        for _ in range(3):
            copies()
        copies_join()
        barrier()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record()
        for _ in range(steps):
            copies()
        copies_join()
        c1.record()
        barrier()
        tc = torch.tensor([c0.elapsed_time(c1)], device=dev)
        dist.all_reduce(tc, op=dist.ReduceOp.MAX)
        e2e["copies_only_ms_per_step"] = float(tc.item()) / steps
        e2e["copies_only_gbs_per_gpu"] = (e2e["h2d_bytes_per_step"] + e2e["d2h_bytes_per_step"]) / (float(tc.item()) / steps * 1e-3) / 1e9

    # dominant kernel alone: every GroupGEMM launch of the step on this rank (routed lists + shared-expert lists)
    gemm = lambda: layer.gemm_only()
    for _ in range(3):
        gemm()
    ms_gemm = timed(gemm, steps)
    pk = peaks()
    int8_peak = 2.0 * pk["bf16"]
    counts = blocks[-1].last_counts()
    if wl["q"] in ("fp16",) or "a16" in wl["q"]:
        # HBM-bound configs: algorithmic bytes = packed weights (+scales) of the experts that received tokens, read once,
        # + activation rows read + fp16 outputs written by the two launches
        act_bytes = 2 if (wl["q"] == "fp16" or "a16" in wl["q"]) else 1
        rows = sum(counts)
        w_bytes = blocks[-1].weight_bytes(active=[c > 0 for c in counts])
        io_bytes = rows * (s["H"] * act_bytes + 2 * s["I"] * 2 + s["I"] * act_bytes + s["H"] * 2)
        if s["ns"]:
            io_bytes += T * (s["H"] * act_bytes + 3 * s["I"] * s["ns"] * 2 + s["H"] * 2)
        ach = (w_bytes + io_bytes) * steps / (ms_gemm * 1e-3) / 1e9
        # dram read + write of the two launches from one `ncu --set full` capture (profiles/ncu_w4a16_stream_r01.txt)
        traffic = 320.7e6 if (args.workload == "c3" and W == 1 and not args.bs) else None
        stream = getattr(blocks[-1], "stream_fused", False)
        roof = {"bound": "hbm", "achieved": ach, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": ach / pk["hbm_gbs"], "traffic": traffic,
                "kernel": ("gg_pair_kernel" if blocks[-1].fused_silu else "w4a16_stream_kernel" if stream else
                           "gg_persistent_kernel + w4a16_stream_kernel") + " (gate_up + down launches of one step)",
                "peak_source": pk["src"] + " copy bandwidth",
                "algorithmic": f"{w_bytes} B packed weights of the active experts + {io_bytes} B activations in/out per step",
                "ms_per_launch_pair": ms_gemm / steps}
    else:
        ach = my_gemm_ops * steps / (ms_gemm * 1e-3) / 1e12
        # dram__bytes_read.sum + dram__bytes_write.sum of the two launches from one `ncu --set full` capture of this command
        # (profiles/ncu_gg_pair_r02.txt: 315.4 + 139.1 MB gate_up, 212.9 + 149.7 MB down); only known for the default
        # workload on one GPU
        traffic = 817.1e6 if (args.workload == "c2" and W == 1 and not args.bs) else None
        issue_peak = {"int8": 4590.0, "f16": 2232.0}
        roof = {"bound": "tensor", "achieved": ach, "peak": int8_peak, "unit": "TOPS", "frac": ach / int8_peak, "traffic": traffic,
                "kernel": ("gg_pair_kernel" if blocks[-1].fused_silu else "gg_persistent_kernel") +
                          f" ({2 * nL} launches/step on the routed lists" + (" + 2 on the shared-expert lists" if (W > 1 and s["ns"]) else "") +
                          ": gate_up with fused silu*mul epilogue, down)",
                "peak_source": f"2 x {pk['src']} cuBLAS bf16 burst ({pk['bf16']} TFLOPS): int8 dense = 2x bf16 on sm_100",
                "algorithmic": "sum 2*M*N*K of the problems of the timed launches", "ms_per_launch_pair": ms_gemm / steps,
                # second denominator: what one SM-pair's MMA issue reaches with operands resident in shared memory
                # (tools/mma_peak.cu on this pool, profiles/mma_peak_r01.log: 4590 TOPS kind::i8, 2232 TFLOPS kind::f16)
                "frac_of_mma_issue_peak": ach / issue_peak["int8"], "mma_issue_peak": issue_peak["int8"],
                "tensor_pipe_active_ncu": NCU_TENSOR_PIPE.get(args.workload) if (W == 1 and not args.bs) else None}
        if sustained is not None:
            # the GEMM launches alone looped >= 2 s, against the sustained cuBLAS figure (seconds-long loop under the power cap)
            n_g = max(steps, int(2000.0 / max(ms_gemm / steps, 1e-3)) + 1)
            smp3 = ClockSampler(local).begin()
            ms_g = timed(gemm, n_g)
            ck3 = smp3.finish()
            ach_s = my_gemm_ops * n_g / (ms_g * 1e-3) / 1e12
            roof["sustained"] = {"achieved": ach_s, "peak": 2.0 * pk["bf16_sustained"], "frac": ach_s / (2.0 * pk["bf16_sustained"]),
                                 "peak_source": f"2 x {pk['src']} cuBLAS bf16 sustained ({pk['bf16_sustained']} TFLOPS)",
                                 "seconds": ms_g * 1e-3, "clocks": ck3}
    if sustained is not None and roof["bound"] == "tensor":
        sustained["frac_of_2x_bf16_sustained"] = sustained["value"] / W / (2.0 * pk["bf16_sustained"])

    line = {"metric": "MoE-layer GroupGEMM TOPS", "value": value, "unit": "TOPS", "n_gpus": W, "steps": steps, "warmup": warmup,
            "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": ("fp8 (E4M3) tensor cores, fp32 accumulate, fp16 out" if "E4M3" in wl["q"] else
                      "int8 tensor cores over int4/int8 grids, fp16 out") if roof["bound"] == "tensor" else "fp16", "data": "synthetic",
            "config": cfg, "clocks": clocks, "e2e": e2e, "gpu_launches": layer.launches_per_step() * steps,
            "roofline": roof}
    if sustained is not None:
        line["sustained"] = sustained
    if parity is not None:
        line["parity_check"] = parity
    if per_block is not None:
        line["per_block_ms"] = per_block
        line["config"]["blocks"] = nL
    rc = 0
    if parity is not None and not parity["bit_identical"]:
        rc = 3
    if rank == 0:
        if not args.no_cpu and W == 1:
            line["cpu_baseline"] = run_cpu(wl, 3, 3)
        if W == 1 and not args.bs and args.workload in GPU_REFERENCE:
            line["gpu_reference"] = GPU_REFERENCE[args.workload]
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    if rc:
        sys.exit(rc)


# The reference's OWN GroupGEMM kernel on the same B200 (its cta_gemm templates compiled for sm_100a, oracle/ref_gg_harness.cu),
# measured by tools/ref_kernel_compare.py next to mxb_group_gemm on the same problem lists — a RECORD of that run: bench.py
# itself never loads anything under oracle/ outside its CPU legs.
GPU_REFERENCE = {"c2": {"unit": "TOPS", "gate_up": {"reference_kernel": 46.1, "mxb_group_gemm": 2384.0},
                        "down": {"reference_kernel": 43.7, "mxb_group_gemm": 2197.4}, "speedup": 51,
                        "kernel": "cta_gemm_multistage_v2 / _qab_v2, 128 x 128 tiles, mma.sync (s4 emulated on sm_100), grid = #SM",
                        "source": "profiles/ref_kernel_compare_r02.jsonl (tools/ref_kernel_compare.py)"}}

# sm__pipe_tensor_cycles_active (% of elapsed) of the GroupGEMM launches from the committed ncu captures of the default command
# (round 2: token chunks of >= 160 rows run with UMMA M = 256 over the tokens, so this figure now includes the tensor time of
# rows that do not exist — 192 of 256 for c2's routed experts; `achieved` counts algorithmic operations only)
NCU_TENSOR_PIPE = {"c2": {"gate_up": 0.68, "down": 0.65, "source": "profiles/ncu_gg_pair_r02.txt",
                          "note": "includes padded token rows of token-major tiles; round 1 (all weight-major): 0.54 / 0.57"}}


if __name__ == "__main__":
    main()
