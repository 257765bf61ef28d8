"""Workload (problem list) generation — same JSON schema and arithmetic as the reference's
mxmoe/kernels/gen_workload.py:38-110, plus synthetic gate traces in the moe-gate.json schema
(mxmoe/quant/moe_tracer.py:174-211) because the reference's calibration traces are not shipped.
"""
from __future__ import annotations

import json
import os
from typing import Dict, List, Optional

import numpy as np

from .config import MODEL_SHAPES, QLinearConfig


def freq_to_prob(freq: List[int]) -> List[float]:
    total = sum(freq)
    return [f / total for f in freq]


def synthetic_gate_trace(model: str, layers: List[int], kind: str = "uniform", seed: int = 0, num_tokens: int = 4096) -> dict:
    """A gate trace with the keys gen_workload reads: topk, NK=[I,H], num_shared_experts, layer-i.access_freq.
    kind: 'uniform' | 'zipf' (s=1.0) | 'fixture' (the M distribution of the reference's shipped
    out/workloads/qwen2_moe-wiki2-8192-fp16.json, layer 10)."""
    s = MODEL_SHAPES[model]
    rng = np.random.default_rng(seed)
    tr = {"topk": s["topk"], "NK": [s["I"], s["H"]], "num_layers": len(layers), "num_tokens": num_tokens, "num_samples": 1,
          "num_shared_experts": s["ns"]}
    for L in layers:
        E = s["E"]
        if kind == "uniform":
            freq = [num_tokens * s["topk"] // E] * E
        elif kind == "zipf":
            w = 1.0 / np.arange(1, E + 1)
            w = rng.permutation(w / w.sum())
            freq = np.maximum(1, np.round(w * num_tokens * s["topk"])).astype(int).tolist()
        elif kind == "fixture":
            fx = json.load(open(os.path.join(os.path.dirname(__file__), "data", "fixture_qwen2_moe_8192.json")))
            freq = [p[0] for p in fx["gate_up"][:E]]
        else:
            raise ValueError(kind)
        tr[f"layer-{L}"] = {"access_freq": [int(f) for f in freq]}
    return tr


def generate_workload_from_gate_trace(trace: dict, num_total_tokens: int, layer_id: int, qconfig: Optional[dict] = None,
                                      qstr: Optional[str] = None) -> dict:
    """gen_workload.py:38-110 on in-memory dicts.  Returns {"num_tokens": T, "layer-L": {"gate_up": [...], "down": [...]}}
    with entries {"shape": [M,N,K], "w_bits", "a_bits", "gsize", "sym"}.  Like the reference, the shared expert's
    *down* problem takes the *gate* config (gen_workload.py:101)."""
    assert qconfig is None or qstr is None
    if qconfig is not None:
        qconfig = {k: v for k, v in qconfig.items() if k != "LT"}
        qcfg = None
    else:
        q = QLinearConfig.from_qstr(qstr)
        qcfg = {"w_bits": q.w_bits, "a_bits": q.a_bits, "gsize": q.w_gsize, "sym": q.w_sym if qstr is not None else True}
    topk = trace["topk"]
    N, K = trace["NK"]
    ns = trace["num_shared_experts"]
    result = {"num_tokens": num_total_tokens}
    for key, value in trace.items():
        if not key.startswith("layer-"):
            continue
        layer_idx = key.split("-")[1]
        if layer_id != -1 and layer_idx != str(layer_id):
            continue

        def lin(exp_idx: str, linear: str) -> dict:
            if qconfig is None:
                return qcfg
            c = qconfig[layer_idx]["experts"][exp_idx][linear]
            return {"w_bits": c["w_bits"], "a_bits": c["a_bits"], "gsize": c["w_gsize"], "sym": c["w_sym"]}

        num_experts = len(value["access_freq"])
        prob = freq_to_prob(value["access_freq"])
        shapes = {"gate_up": [], "down": []}
        for e, f in enumerate(prob):
            m = int(f * num_total_tokens * topk)
            shapes["gate_up"].append({"shape": [m, N * 2, K], **lin(str(e), "gate")})
            shapes["down"].append({"shape": [m, K, N], **lin(str(e), "down")})
        if ns != 0:
            shapes["gate_up"].append({"shape": [int(num_total_tokens), N * 2 * ns, K], **lin(str(num_experts), "gate")})
            shapes["down"].append({"shape": [int(num_total_tokens), K, N * ns], **lin(str(num_experts), "gate")})
        result[f"layer-{layer_idx}"] = shapes
    return result


def save_workload(workload: dict, path: str) -> None:
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "w") as f:
        json.dump(workload, f, indent=2)


def load_workload(path: str) -> dict:
    with open(path) as f:
        return json.load(f)


def workload_ops(workload: dict) -> int:
    """sum of 2*M*N*K over every problem (the reference's TFLOPS numerator, test.cu:109-116)."""
    tot = 0
    for k, v in workload.items():
        if not k.startswith("layer-"):
            continue
        for part in ("gate_up", "down"):
            for p in v[part]:
                m, n, kk = p["shape"]
                tot += 2 * m * n * kk
    return tot


def synthetic_mixed_qconfig(model: str, layers: List[int], wbits: float = 5.0, seed: int = 0) -> dict:
    """An ILP-like assignment in the bits_solver.export_qconfig schema (bits_solver.py:42-53): each expert's
    gate/up (tied, bits_solver.py:378-379) and down blocks are w8a8_g-1_sym or w4a4_g-1_sym such that the
    size-weighted average weight width is `wbits` (5.0 -> a quarter of the blocks are 8-bit)."""
    s = MODEL_SHAPES[model]
    rng = np.random.default_rng(seed)
    E = s["E"]

    def lin(b):
        return {"w_bits": b, "w_gsize": -1, "w_sym": True, "w_clip": [1.0, 1.0], "a_bits": b, "a_gsize": -1, "a_sym": True,
                "a_clip": [1.0, 1.0]}

    frac8 = (wbits - 4.0) / 4.0
    out: Dict[str, dict] = {}
    for L in layers:
        units = E + (1 if s["ns"] else 0)
        # block list: (expert, 'gu' weight 2 | 'down' weight 1), shared expert scaled by ns
        blocks = []
        for e in range(units):
            scale = s["ns"] if e == E else 1
            blocks.append((e, "gu", 2 * scale))
            blocks.append((e, "down", 1 * scale))
        order = rng.permutation(len(blocks))
        total = sum(b[2] for b in blocks)
        budget8 = frac8 * total
        choice = {}
        acc = 0.0
        for i in order:
            e, kind, wt = blocks[i]
            if acc + wt <= budget8 + 1e-9:
                choice[(e, kind)] = 8
                acc += wt
            else:
                choice[(e, kind)] = 4
        ex = {}
        for e in range(units):
            gu = lin(choice[(e, "gu")])
            ex[str(e)] = {"gate": gu, "up": dict(gu), "down": lin(choice[(e, "down")])}
        out[str(L)] = {"experts": ex}
    out["LT"] = {str(L): [0.0, 0.0] for L in layers}
    return out
