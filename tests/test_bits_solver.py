"""Bit-width allocation (mxmoe_b200/bits_solver.py) against the reference's own ILP instance.

tests/golden/bits_model_1.json is the model the reference's solver wrote for layer 1 of Qwen1.5-MoE
(bits_model-1.lp, bits_solver.py:384; extracted by tests/golden/gen_bits_golden.py): 61 units x 3 blocks x 2 strategies
(4 and 8 bit), budget 960, r = 1.  The tests check that our constraint builder produces THAT model (same coefficients,
same number of constraints of every family) and that the HiGHS optimum equals an independent exact solver
(a multiple-choice-knapsack DP over the experts)."""
import json
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "bits_model_1.json")


@pytest.fixture(scope="module")
def inst():
    return json.load(open(GOLD))


def dp_optimum(delta, mem, budget):
    """Exact optimum by dynamic programming: per expert 4 choices (gate = up tied, down free), integer weights."""
    delta, mem = np.asarray(delta), np.asarray(mem)
    E, _, S = delta.shape
    B = int(budget)
    best = np.full(B + 1, np.inf)
    best[0] = 0.0
    for e in range(E):
        nxt = np.full(B + 1, np.inf)
        for sg in range(S):
            for sd in range(S):
                w = int(mem[e, 0, sg] + mem[e, 1, sg] + mem[e, 2, sd])
                c = delta[e, 0, sg] + delta[e, 1, sg] + delta[e, 2, sd]
                if w <= B:
                    nxt[w:] = np.minimum(nxt[w:], best[:B + 1 - w] + c)
        best = nxt
    return float(best.min())


def test_constraint_builder_reproduces_the_reference_model(inst):
    from mxmoe_b200.bits_solver import build_constraints
    E, S = inst["E"], inst["S"]
    A_ub, b_ub, A_eq, b_eq = build_constraints(E, S, np.asarray(inst["mem_coef"]), inst["budget"])
    assert A_ub.shape == (1, E * 3 * S) and b_ub[0] == 960.0
    assert np.array_equal(A_ub.reshape(E, 3, S), np.asarray(inst["mem_coef"]))          # memory_budget row
    assert int((b_eq == 1).sum()) == inst["n_one_strategy"] == E * 3                     # 1_strategy_for_E*_W*
    assert int((b_eq == 0).sum()) == inst["n_gate_eq_up"] == E * S                       # x_equal_for_E*_S*
    one = A_eq[:E * 3]
    assert np.all(one.sum(1) == S) and np.all((one == 0) | (one == 1))
    eq = A_eq[E * 3:]
    assert np.all(eq.sum(1) == 0) and np.all(np.abs(eq).sum(1) == 2)


def test_highs_optimum_equals_exact_dp_on_the_reference_instance(inst):
    from mxmoe_b200.bits_solver import solve_layer
    assign, L, T = solve_layer(inst["delta"], inst["mem_coef"], inst["budget"], None, r=1.0)
    a = np.asarray(assign)
    assert a.shape == (61, 3) and np.array_equal(a[:, 0], a[:, 1])                      # gate == up
    mem = np.asarray(inst["mem_coef"])
    used = sum(mem[e, n, a[e, n]] for e in range(61) for n in range(3))
    assert used <= inst["budget"]
    assert L == pytest.approx(dp_optimum(inst["delta"], inst["mem_coef"], inst["budget"]), rel=1e-9)
    # the budget (5.0 bits average) binds: the all-8-bit assignment is infeasible, so some blocks drop to 4 bit
    assert 0 < int((a == 0).sum()) < a.size


def test_tradeoff_moves_along_the_pareto_front():
    """r sweeps accuracy-only -> runtime-only on the B200 cost model (mxb_select_tiles through the C-ABI, no GPU needed):
    loss is non-decreasing and estimated runtime non-increasing; the exported qconfig round-trips through the reference
    schema and hits the bit budget."""
    from mxmoe_b200 import config as C
    from mxmoe_b200.bits_solver import average_wbits, solve_model
    prev = None
    for r in (1.0, 0.5, 0.0):
        q = solve_model("qwen2_moe", ["w4a4_g-1_sym", "w8a8_g-1_sym"], 5.0, 8192, r=r, layers=[11])
        L, T = q["LT"]["11"]
        assert average_wbits(q, "qwen2_moe") <= 5.0 + 1e-9
        ql = C.qmodel_cfg_from_dict({k: v for k, v in q.items() if k != "LT"}).layers["11"]
        assert len(ql.experts) == 61
        for e in ql.experts.values():
            assert (e.gate.w_bits, e.gate.a_bits) == (e.up.w_bits, e.up.a_bits)
        if prev is not None and r < 1.0:
            assert L >= prev[0] - 1e-9
        prev = (L, T)
    q1 = solve_model("qwen2_moe", ["w4a4_g-1_sym", "w8a8_g-1_sym"], 5.0, 8192, r=0.5, layers=[11])
    q0 = solve_model("qwen2_moe", ["w4a4_g-1_sym", "w8a8_g-1_sym"], 5.0, 8192, r=0.0, layers=[11])
    assert q0["LT"]["11"][1] <= q1["LT"]["11"][1] + 1e-9


def test_measured_cost_table_drives_the_solver():
    """profiles/cost_table_b200_r01.json (tools/cost_table.py on B200: measured microseconds per expert problem and
    strategy) as the runtime model, the role of the reference's perf/performance_table.json: log-log interpolation in the
    row count is monotone, and the runtime-only optimum (r = 0) prefers the measured-faster strategy wherever the budget
    allows."""
    from mxmoe_b200 import workload as Wl
    from mxmoe_b200.bits_solver import cost_from_table, solve_model
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    table = json.load(open(os.path.join(root, "profiles", "cost_table_b200_r01.json")))
    trace = Wl.synthetic_gate_trace("qwen2_moe", [11], "fixture")
    lw = Wl.generate_workload_from_gate_trace(trace, 8192, 11)["layer-11"]
    strategies = ["w4a4_g-1_sym", "w8a8_g-1_sym"]
    cost = cost_from_table(table, lw, strategies)
    assert cost.shape == (61, 3, 2) and (cost > 0).all()
    rows = np.array([p["shape"][0] for p in lw["gate_up"][:60]])
    order = np.argsort(rows)
    assert np.all(np.diff(cost[:60, 0, 1][order]) >= -1e-9)       # more rows never cost less
    assert cost_from_table(table, lw, ["w2a16_g128_asym"]) is None  # unknown strategy -> caller falls back to the selector
    q = solve_model("qwen2_moe", strategies, 8.0, 8192, r=0.0, layers=[11], cost_table=table)
    picked8 = sum(1 for e in q["11"]["experts"].values() if e["down"]["w_bits"] == 8)
    faster8 = int((cost[:, 2, 1] < cost[:, 2, 0]).sum())
    assert picked8 >= faster8  # with a budget of 8 bits nothing stops the faster choice
