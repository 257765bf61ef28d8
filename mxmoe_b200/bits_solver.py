"""Bit-width allocation for the MoE blocks on a B200 cost model: the co-design loop of the reference
(mxmoe/quant/bits_solver.py) rebuilt on scipy's HiGHS MILP and on the runtime estimates of THIS library's tile selector.

What the reference does (bits_solver.py:201-422): per layer, binary x[e,n,s] = linear block n (gate, up, down) of expert e
takes quantization strategy s (a qstr such as `w4a4_g-1_sym`); minimise  L^r * T^(1-r)  with
  L = sum x * delta[e][n][s]        (calibration loss of the block under the strategy, get_strategy_loss :85-101)
  T = sum x * cost[e][n][s]         (runtime of the block's GEMM problem, get_runtime_cost :518-542)
subject to the memory budget  sum wbits[s] * size[e] * x <= wbits_target * sum sizes  (:361-366), one strategy per block
(:368-369) and gate == up (:378-379, they are one fused GEMM).  It needs Gurobi (absent here) and a per-tile cost table
measured on an RTX 4090 (perf/performance_table.json, keyed by tile-config reprs that do not exist on Blackwell).

Here:
  * the solver is `scipy.optimize.milp` (HiGHS).  r = 1 and r = 0 are linear; for 0 < r < 1 the product is minimised over
    the Pareto front by an epsilon-constraint sweep (min L s.t. T <= tau for a grid of tau) — every optimum of
    L^r * T^(1-r) is Pareto-optimal, the grid bounds the gap;
  * cost[e][n][s] comes from `mxb_select_tiles` (include/mxmoe_b200.h): the kernel and tile this library would run the
    problem with and its roofline time on B200 — the tile dimension t of the reference's y[e,n,s,t] disappears because
    the selector fixes the tile per problem (there is no cross-problem "equal warp count" fusion constraint any more,
    compose_kernel.py:69-71);
  * the output is the reference's qconfig JSON (export_qconfig :25-71, read by qconfig.py:66-80 and run_mxmoe_gg.py:18-28),
    so `run_mxmoe_gg.py --qconfig` consumes it unchanged.

CLI:  python -m mxmoe_b200.bits_solver --model qwen2_moe --wbits 5.0 --batch 8192 --r 0.75 \
          --filter_list w4a4_g-1_sym w8a8_g-1_sym [--loss loss.json] [--out qconfig.json]
Without --loss a synthetic loss profile is used (seeded; larger for popular experts and for fewer bits) so that the
pipeline can be exercised without the accuracy side (calibration is out of scope).
"""
from __future__ import annotations

import argparse
import json
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import config as C
from . import workload as W

BLOCKS = ("gate", "up", "down")


# ------------------------------------------------------------------------------------------------ the MILP
def build_constraints(E: int, S: int, mem_coef: np.ndarray, budget: float):
    """Constraint matrices over x flattened as [(e * 3 + n) * S + s] (bits_solver.py:361-379).

    Returns (A_ub, b_ub, A_eq, b_eq): memory budget; one strategy per block; gate == up per (expert, strategy)."""
    nv = E * 3 * S
    idx = lambda e, n, s: (e * 3 + n) * S + s
    A_ub = np.asarray(mem_coef, dtype=np.float64).reshape(1, nv)
    b_ub = np.array([float(budget)])
    rows, rhs = [], []
    for e in range(E):
        for n in range(3):
            r = np.zeros(nv)
            r[[idx(e, n, s) for s in range(S)]] = 1.0
            rows.append(r)
            rhs.append(1.0)
    for e in range(E):
        for s in range(S):
            r = np.zeros(nv)
            r[idx(e, 0, s)] = 1.0
            r[idx(e, 1, s)] = -1.0
            rows.append(r)
            rhs.append(0.0)
    return A_ub, b_ub, np.array(rows), np.array(rhs)


def _milp(c, A_ub, b_ub, A_eq, b_eq, extra: Optional[Tuple[np.ndarray, float]] = None):
    from scipy.optimize import Bounds, LinearConstraint, milp
    cons = [LinearConstraint(A_ub, -np.inf, b_ub), LinearConstraint(A_eq, b_eq, b_eq)]
    if extra is not None:
        cons.append(LinearConstraint(extra[0].reshape(1, -1), -np.inf, extra[1]))
    res = milp(c=c, constraints=cons, integrality=np.ones(len(c)), bounds=Bounds(0, 1))
    if not res.success:
        return None
    return np.rint(res.x).astype(np.int64)


def solve_layer(delta, mem_coef, budget: float, cost=None, r: float = 1.0, n_eps: int = 16):
    """delta, mem_coef, cost: [E][3][S].  Returns (assign [E][3] strategy index, L, T).

    r = 1: minimise the loss (the instance of bits_model-1.lp); r = 0: the runtime; else L^r * T^(1-r)."""
    delta = np.asarray(delta, dtype=np.float64)
    E, _, S = delta.shape
    cost = np.zeros_like(delta) if cost is None else np.asarray(cost, dtype=np.float64)
    A_ub, b_ub, A_eq, b_eq = build_constraints(E, S, mem_coef, budget)
    d, t = delta.reshape(-1), cost.reshape(-1)

    def measure(x):
        return float(d @ x), float(t @ x)

    if r >= 1.0 or not cost.any():
        x = _milp(d, A_ub, b_ub, A_eq, b_eq)
    elif r <= 0.0:
        x = _milp(t, A_ub, b_ub, A_eq, b_eq)
    else:
        x_loss = _milp(d, A_ub, b_ub, A_eq, b_eq)       # slowest point of the front
        x_fast = _milp(t, A_ub, b_ub, A_eq, b_eq)       # fastest
        if x_loss is None or x_fast is None:
            raise ValueError("infeasible: the memory budget is below the smallest strategy")
        best, x = None, None
        t_lo, t_hi = measure(x_fast)[1], measure(x_loss)[1]
        for tau in np.linspace(t_lo, t_hi, max(2, n_eps)):
            xi = _milp(d, A_ub, b_ub, A_eq, b_eq, extra=(t, tau * (1 + 1e-9)))
            if xi is None:
                continue
            L, T = measure(xi)
            val = r * np.log(max(L, 1e-30)) + (1 - r) * np.log(max(T, 1e-30))
            if best is None or val < best:
                best, x = val, xi
    if x is None:
        raise ValueError("infeasible: the memory budget is below the smallest strategy")
    L, T = measure(x)
    assign = x.reshape(E, 3, S).argmax(axis=2)
    return assign.tolist(), L, T


# ------------------------------------------------------------------------------------------------ B200 runtime cost
def strategy_bits(qstr: str) -> Tuple[int, int]:
    """(w_bits, a_bits) as counted against the budget (bits_solver.py:581-596); E4M3 counts 8."""
    q = C.QLinearConfig.from_qstr(qstr)
    return abs(q.w_bits), abs(q.a_bits)


def cost_from_table(table: dict, layer_workload: Dict[str, list], strategies: Sequence[str]) -> Optional[np.ndarray]:
    """cost[e][n][s] from a MEASURED table (tools/cost_table.py -> profiles/cost_table_b200_r01.json: per strategy and block,
    microseconds of one expert problem at a ladder of row counts), log-log interpolated in the row count.  The table plays
    the part of the reference's perf/performance_table.json (bits_solver.py:533-541).  None if a strategy is missing."""
    gu, dn = layer_workload["gate_up"], layer_workload["down"]
    E = len(gu)
    cost = np.zeros((E, 3, len(strategies)))
    for s, qs in enumerate(strategies):
        ent = table.get("strategies", {}).get(qs)
        if ent is None:
            return None
        for blk, lst, cols in (("gate_up", gu, (0, 1)), ("down", dn, (2,))):
            pts = sorted((int(k.split(":")[1]), v["measured_us"], v["shape"]) for k, v in ent.items() if k.startswith(blk + ":"))
            ms, us = np.log([p[0] for p in pts]), np.log([p[1] for p in pts])
            n_ref = pts[0][2][1]
            for e, p in enumerate(lst):
                M, N, _ = p["shape"]
                t = float(np.exp(np.interp(np.log(max(M, 1)), ms, us))) * (N / n_ref)  # wider experts (shared): linear in N
                for c in cols:
                    cost[e, c, s] = t / len(cols)
    return cost


def runtime_cost(model: str, layer_workload: Dict[str, list], strategies: Sequence[str], handle=None) -> np.ndarray:
    """cost[e][n][s] in microseconds from the library's own selector (mxb_select_tiles: kernel path, tile and roofline
    time per problem).  gate and up are one fused GEMM (N = 2I): its estimate is split evenly between the two blocks,
    which the gate == up constraint keeps together anyway (ref bits_solver.py:545-579 builds the same shapes)."""
    import ctypes

    from . import ops
    lib = ops.load_library()
    gu, dn = layer_workload["gate_up"], layer_workload["down"]
    E = len(gu)
    cost = np.zeros((E, 3, len(strategies)))
    for s, qs in enumerate(strategies):
        q = C.QLinearConfig.from_qstr(qs)
        probs = []
        for lst in (gu, dn):
            for p in lst:
                M, N, K = p["shape"]
                probs.append(ops.MxbProblem(None, None, None, None, None, None, 0, 0, max(int(M), 1), int(N), int(K), q.a_bits, q.w_bits,
                                            q.w_gsize, 1 if (q.w_sym or q.w_bits >= 16) else 0, 0, 0, None, None, 0))
        arr = (ops.MxbProblem * len(probs))(*probs)
        out = (ops.MxbTileChoice * len(probs))()
        rc = lib.mxb_select_tiles(handle.h if handle is not None else None, arr, len(probs), out)
        if rc != 0:
            raise ops.MxbError(f"mxb_select_tiles failed for {qs}: {lib.mxb_last_error().decode()}")
        for e in range(E):
            cost[e, 0, s] = cost[e, 1, s] = 0.5 * out[e].est_us
            cost[e, 2, s] = out[E + e].est_us
    return cost


# ------------------------------------------------------------------------------------------------ model level
def synthetic_loss(E: int, strategies: Sequence[str], access_prob: Sequence[float], seed: int = 0) -> np.ndarray:
    """Stand-in for the calibration losses (bits_solver.py:85-101 reads them from the accuracy pipeline's JSON): grows as
    4^(-bits) like uniform quantization noise, scaled by how often the expert is used, log-normal jitter per block."""
    rng = np.random.default_rng(seed)
    d = np.zeros((E, 3, len(strategies)))
    p = np.asarray(access_prob, dtype=np.float64)
    p = p / p.mean()
    for s, qs in enumerate(strategies):
        w, a = strategy_bits(qs)
        base = 4.0 ** (8 - w) + (0.25 * 4.0 ** (8 - a) if a < 16 else 0.0)
        d[:, :, s] = base * p[:, None] * rng.lognormal(0.0, 0.3, size=(E, 3))
    return d


def solve_model(model: str, strategies: Sequence[str], wbits: float, batch: int, r: float = 1.0, layers: Optional[Sequence[int]] = None,
                loss: Optional[Dict[str, np.ndarray]] = None, trace_kind: str = "fixture", seed: int = 0,
                cost_table: Optional[dict] = None):
    """Layer-by-layer allocation (solve_model_qconfig_layer_level).  Returns (qconfig dict in the reference schema, LT)."""
    s = C.MODEL_SHAPES[model]
    layers = list(layers) if layers is not None else [0]
    trace = W.synthetic_gate_trace(model, layers, trace_kind if model == "qwen2_moe" else "uniform")
    units = s["E"] + (1 if s["ns"] else 0)
    sizes = [1.0] * s["E"] + ([float(s["ns"])] if s["ns"] else [])  # relative weight sizes (bits_solver.py:636-644)
    sbits = [strategy_bits(q)[0] for q in strategies]
    mem = np.array([[[sbits[k] * sizes[e] for k in range(len(strategies))] for _ in range(3)] for e in range(units)])
    budget = wbits * 3 * sum(sizes)
    out: Dict[str, dict] = {"LT": {}}
    for li in layers:
        key = f"layer-{li}"
        freq = list(trace[key]["access_freq"]) + ([float(np.mean(trace[key]["access_freq"])) * s["topk"]] if s["ns"] else [])
        d = loss[str(li)] if loss is not None else synthetic_loss(units, strategies, freq, seed + li)
        cost = None
        if r < 1.0:
            lw = W.generate_workload_from_gate_trace(trace, batch, li)[key]
            cost = cost_from_table(cost_table, lw, strategies) if cost_table is not None else None
            if cost is None:  # no measured table (or a strategy it lacks): the selector's roofline estimate
                cost = runtime_cost(model, lw, strategies)
        assign, L, T = solve_layer(d, mem, budget, cost, r)
        experts = {}
        for e in range(units):
            experts[str(e)] = {b: C.QLinearConfig.from_qstr(strategies[assign[e][n]]).to_dict() for n, b in enumerate(BLOCKS)}
        out[str(li)] = {"experts": experts}
        out["LT"][str(li)] = [L, T]
    return out


def tile_config_for(model: str, qcfg: dict, batch: int, trace_kind: str = "fixture") -> Dict[str, str]:
    """The `_tile_cfg.json` the reference writes next to the qconfig (export_qconfig, bits_solver.py:30,67-68):
    {layer: repr of one TileConfig per strategy used in the layer}.  Here the tile comes from `mxb_select_tiles` for the
    layer's mean routed gate_up problem under that strategy; PATH / QCFG are this library's additions to the repr
    (config.tile_for / tile_hint read them back; run_mxmoe_gg.py --tile_config consumes the file)."""
    import ctypes  # noqa: F401

    from . import ops
    lib = ops.load_library()
    s = C.MODEL_SHAPES[model]
    M = max(1, batch * s["topk"] // s["E"])
    out = {}
    for li, layer in qcfg.items():
        if li == "LT":
            continue
        used = sorted({C.QLinearConfig.from_dict(b[n]).qstr() for b in layer["experts"].values() for n in BLOCKS})
        reps = []
        for qs in used:
            q = C.QLinearConfig.from_qstr(qs)
            pr = ops.MxbProblem(None, None, None, None, None, None, 0, 0, M, 2 * s["I"], s["H"], q.a_bits, q.w_bits, q.w_gsize,
                                1 if (q.w_sym or q.w_bits >= 16) else 0, 0, 0, None, None, 0)
            arr = (ops.MxbProblem * 1)(pr)
            ch = (ops.MxbTileChoice * 1)()
            if lib.mxb_select_tiles(None, arr, 1, ch) != 0:
                raise ops.MxbError(f"mxb_select_tiles failed for {qs}: {lib.mxb_last_error().decode()}")
            t = ch[0]
            mma = ("UMMA_F16" if q.a_bits >= 16 else "UMMA_F8F6F4" if q.w_bits == C.BITS_E4M3 else "UMMA_I8") + ("_2CTA" if t.path == 2 else "")
            reps.append(f"TileConfig(BM={t.tile_m}, BN={t.tile_n}, BK={t.tile_k}, WM=1, WN=1, WK=1, STAGE={t.stages}, SPLITK=-1, "
                        f"MMA='{mma}', PATH={t.path}, QCFG='{qs}')")
        out[str(li)] = "(" + ", ".join(reps) + ("," if len(reps) == 1 else "") + ")"
    return out


def average_wbits(qcfg: dict, model: str) -> float:
    s = C.MODEL_SHAPES[model]
    sizes = [1.0] * s["E"] + ([float(s["ns"])] if s["ns"] else [])
    tot = num = 0.0
    for li, layer in qcfg.items():
        if li == "LT":
            continue
        for e, blocks in layer["experts"].items():
            for b in BLOCKS:
                tot += abs(blocks[b]["w_bits"]) * sizes[int(e)]
                num += sizes[int(e)]
    return tot / num


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--model", default="qwen2_moe", choices=sorted(C.MODEL_SHAPES))
    ap.add_argument("--wbits", type=float, default=5.0)
    ap.add_argument("--batch", type=int, default=8192)
    ap.add_argument("--r", type=float, default=1.0, help="1 = accuracy only, 0 = runtime only (bits_solver.py --r)")
    ap.add_argument("--layers", type=int, nargs="*", default=[11])
    ap.add_argument("--filter_list", nargs="+", default=["w4a4_g-1_sym", "w8a8_g-1_sym"])
    ap.add_argument("--loss", default=None, help="JSON {layer: [E][3][S]} of calibration losses; default synthetic")
    ap.add_argument("--cost_table", default=None, help="measured table from tools/cost_table.py; default: mxb_select_tiles estimates")
    ap.add_argument("--out", default=None)
    a = ap.parse_args(argv)
    loss = None
    if a.loss:
        loss = {k: np.asarray(v) for k, v in json.load(open(a.loss)).items()}
    table = json.load(open(a.cost_table)) if a.cost_table else None
    q = solve_model(a.model, a.filter_list, a.wbits, a.batch, a.r, a.layers, loss, cost_table=table)
    print(f"average weight bits {average_wbits(q, a.model):.3f} (target {a.wbits}); LT {q['LT']}")
    if a.out:
        json.dump(q, open(a.out, "w"))
        print("wrote", a.out)
        tpath = a.out[:-5] + "_tile_cfg.json" if a.out.endswith(".json") else a.out + "_tile_cfg.json"
        json.dump(tile_config_for(a.model, q, a.batch), open(tpath, "w"))
        print("wrote", tpath)


if __name__ == "__main__":
    main()
