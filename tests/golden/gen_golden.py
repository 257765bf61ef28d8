#!/usr/bin/env python
"""Generate golden fixtures from the REFERENCE itself (run in the build container only —
/root/reference does not exist on the GPU box; the fixtures it writes are committed).

    python tests/golden/gen_golden.py

Writes
  mxmoe_b200/data/hadk.npz          sign bits of the reference's K x K Hadamard matrices
                                    (hadamard_utils.py:160-4379) — data the product needs to be
                                    a drop-in for weights rotated offline by the reference
  tests/golden/quant_minmax.npz     quant_minmax / Quantizer.fake_quant (quant.py:40-110) outputs
  tests/golden/hadamard.npz         matmul_hadU (hadamard_utils.py:62-83) outputs, float64
  tests/golden/workload.json        generate_workload_from_gate_trace (gen_workload.py:38-110)
                                    on a synthetic gate trace + qconfig (also committed)
  tests/golden/pack.npz             outputs of oracle/_ref/ref_pack (the reference's packers,
                                    quantize.cuh:299-475, compiled in place)
  tests/golden/fixture_qwen2_moe_8192.json   problem M list of the shipped workload fixture
"""
import json
import os
import subprocess
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.abspath(os.path.join(HERE, "..", ".."))
REF = "/root/reference"

# stubs for modules the reference imports at module scope but that are absent here
stub = types.ModuleType("fast_hadamard_transform")
stub.hadamard_transform = None
sys.modules["fast_hadamard_transform"] = stub
jsb = types.ModuleType("jsbeautifier")


class _Opt:
    indent_size = 2


jsb.default_options = lambda: _Opt()
jsb.beautify = lambda s, o: json.dumps(json.loads(s), indent=2)
sys.modules["jsbeautifier"] = jsb
sys.path.insert(0, REF)

import mxmoe.quant.hadamard_utils as hu  # noqa: E402
from mxmoe.quant.quant import quant_minmax, Quantizer  # noqa: E402
from mxmoe.kernels.gen_workload import generate_workload_from_gate_trace  # noqa: E402


def gen_hadk():
    out = {}
    for k in (12, 20, 28, 36, 40, 52, 60, 108, 140, 156, 172, 176):
        m = getattr(hu, f"get_had{k}")().numpy()
        assert set(np.unique(m)) == {-1.0, 1.0}
        out[f"had{k}"] = np.packbits((m > 0).astype(np.uint8).reshape(-1))
    np.savez_compressed(os.path.join(ROOT, "mxmoe_b200", "data", "hadk.npz"), **out)


def gen_quant():
    g = torch.Generator().manual_seed(1234)
    out = {}
    cases = [(8, -1, True), (4, -1, True), (4, 128, True), (4, 128, False), (4, -1, False),
             (8, 128, False), (8, -1, False), (2, 128, False)]
    x = (torch.rand(24, 512, generator=g) * 2 - 1).half()
    x[3, 17] = 9.5          # an outlier row
    x[5] = x[5] * 1e-3      # a tiny row
    out["x"] = x.numpy()
    for bits, gs, sym in cases:
        q, s, z = quant_minmax(x.clone(), bits, gs, sym)
        key = f"b{bits}_g{gs}_{'sym' if sym else 'asym'}"
        out[key + "_q"] = q.numpy().astype(np.int32)
        out[key + "_s"] = s.numpy()
        out[key + "_z"] = (z.numpy() if torch.is_tensor(z) else np.zeros_like(s.numpy()))
        out[key + "_fq"] = Quantizer(bits, sym, gs).fake_quant(x.clone()).numpy()
    np.savez_compressed(os.path.join(HERE, "quant_minmax.npz"), **out)


def gen_hadamard():
    g = torch.Generator().manual_seed(7)
    out = {}
    for n in (64, 2048, 1408, 2560, 14336 // 8, 5632, 96):
        x = torch.randn(3, n, generator=g, dtype=torch.float64)
        out[f"x{n}"] = x.numpy()
        out[f"y{n}"] = hu.matmul_hadU(x).numpy()
        _, k = hu.get_hadK(n)
        out[f"k{n}"] = np.int32(k)
    np.savez_compressed(os.path.join(HERE, "hadamard.npz"), **out)


def gen_workload():
    rng = np.random.default_rng(0)
    E = 60
    trace = {"topk": 4, "NK": [1408, 2048], "num_layers": 2, "num_tokens": 4096, "num_samples": 1,
             "num_shared_experts": 4}
    for layer in (10, 11):
        freq = rng.integers(200, 900, size=E).tolist()
        trace[f"layer-{layer}"] = {"access_freq": freq}
    tpath = os.path.join(HERE, "gate_trace.json")
    json.dump(trace, open(tpath, "w"))
    # a mixed qconfig in the bits_solver.export_qconfig schema (bits_solver.py:42-53)
    def lin(w, a, g=-1, sym=True):
        return {"w_bits": w, "w_gsize": g, "w_sym": sym, "w_clip": [1.0, 1.0],
                "a_bits": a, "a_gsize": g, "a_sym": True, "a_clip": [1.0, 1.0]}
    qc = {}
    for layer in (10, 11):
        ex = {}
        for e in range(E + 1):
            gu = lin(8, 8) if rng.random() < 0.25 else lin(4, 4)
            dn = lin(8, 8) if rng.random() < 0.25 else lin(4, 4)
            ex[str(e)] = {"gate": gu, "up": gu, "down": dn}
        qc[str(layer)] = {"experts": ex}
    qc["LT"] = {"10": [0.1, 0.2], "11": [0.1, 0.2]}
    qpath = os.path.join(HERE, "qconfig_mixed.json")
    json.dump(qc, open(qpath, "w"))
    outs = {}
    for name, kw in (("fp16", {}), ("qstr", {"qstr": "w4a16_g128_asym"}), ("qcfg", {"qcfg_file": qpath})):
        save = os.path.join("/tmp", f"wl_{name}.json")
        generate_workload_from_gate_trace(tpath, 8192, 11, save, **kw)
        outs[name] = json.load(open(save))
    json.dump(outs, open(os.path.join(HERE, "workload.json"), "w"))
    fx = json.load(open(os.path.join(REF, "out/workloads/qwen2_moe-wiki2-8192-fp16.json")))
    lay = fx["layer-10"]
    json.dump({"num_tokens": fx["num_tokens"], "layer": 10,
               "gate_up": [[int(v) for v in p["shape"]] for p in lay["gate_up"]],
               "down": [[int(v) for v in p["shape"]] for p in lay["down"]]},
              open(os.path.join(HERE, "fixture_qwen2_moe_8192.json"), "w"))


def ref_pack(op, N, K, M, qbits, sym, gsize, payload: bytes) -> bytes:
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_pack")
    hdr = np.array([op, N, K, M, qbits, int(sym), gsize], np.int32).tobytes()
    return subprocess.run([exe], input=hdr + payload, stdout=subprocess.PIPE, check=True).stdout


def gen_pack():
    rng = np.random.default_rng(5)
    out = {}
    N, K, M = 64, 64, 24
    for bits in (8, 4, 2):
        for sym in (True, False):
            hi = (1 << (bits - 1)) - 1
            q = rng.integers(-hi, hi + 1, size=(N, K)) if sym else rng.integers(0, 1 << bits, size=(N, K))
            q = q.astype(np.int32)
            o = np.frombuffer(ref_pack(0, N, K, 0, bits, sym, -1, q.tobytes()), np.uint16)
            key = f"wonly_b{bits}_{'sym' if sym else 'asym'}"
            out[key + "_q"] = q
            out[key + "_p"] = o.reshape(N // (16 // bits), K)
    for bits in (8, 4):
        hi = (1 << (bits - 1)) - 1
        a = rng.integers(-hi, hi + 1, size=(M, K)).astype(np.int32)
        w = rng.integers(-hi, hi + 1, size=(N, K)).astype(np.int32)
        o = np.frombuffer(ref_pack(1, N, K, M, bits, True, -1, a.tobytes() + w.tobytes()), np.uint16)
        pk = 16 // bits
        out[f"wxax_b{bits}_a"] = a
        out[f"wxax_b{bits}_w"] = w
        out[f"wxax_b{bits}_pa"] = o[: M * K // pk].reshape(M, K // pk)
        out[f"wxax_b{bits}_pw"] = o[M * K // pk:].reshape(N, K // pk)
    for sym in (True, False):
        g = 16
        s = rng.random((N, K // g, 1 if sym else 2)).astype(np.float16)
        o = np.frombuffer(ref_pack(2, N, K, 0, 4, sym, g, s.tobytes()), np.float16)
        out[f"scale_{'sym' if sym else 'asym'}_in"] = s
        out[f"scale_{'sym' if sym else 'asym'}_out"] = o.reshape(K // g, N, 1 if sym else 2)
    np.savez_compressed(os.path.join(HERE, "pack.npz"), **out)


if __name__ == "__main__":
    gen_hadk()
    gen_quant()
    gen_hadamard()
    gen_workload()
    gen_pack()
    print("golden fixtures written")
