"""Extracts the ILP instance the reference's bits_solver wrote for layer 1 of Qwen1.5-MoE (/root/reference/bits_model-1.lp,
written at mxmoe/quant/bits_solver.py:384: budget 960 = 5.0 bits x 192 blocks, 61 units x 3 blocks x 2 strategies
{w4a4_g-1_sym, w8a8_g-1_sym}, r = 1) into tests/golden/bits_model_1.json.  Run here (the reference is not on the GPU box)."""
import json
import os
import re

SRC = "/root/reference/bits_model-1.lp"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "bits_model_1.json")

text = open(SRC).read()
obj = text[text.index("Minimize") + len("Minimize"):text.index("Subject To")]
cons = text[text.index("Subject To") + len("Subject To"):text.index("Bounds")]
term = re.compile(r"([+-]?\s*[0-9.eE+-]*)\s*x\[(\d+),(\d+),(\d+)\]")


def terms(s):
    out = {}
    for c, e, n, k in term.findall(s):
        c = c.replace(" ", "")
        v = 1.0 if c in ("", "+") else (-1.0 if c == "-" else float(c))
        out[(int(e), int(n), int(k))] = v
    return out


delta = terms(obj)
E = max(k[0] for k in delta) + 1
N = max(k[1] for k in delta) + 1
S = max(k[2] for k in delta) + 1
# constraints: split on names
blocks = re.split(r"\n\s*([A-Za-z_0-9]+):", "\n" + cons)
names, bodies = blocks[1::2], blocks[2::2]
mem = None
n_one, n_eq, n_y = 0, 0, 0
for nm, body in zip(names, bodies):
    body = " ".join(body.split())
    if nm == "memory_budget":
        lhs, rhs = body.split("<=")
        mem = (terms(lhs), float(rhs))
    elif nm.startswith("1_strategy_for"):
        n_one += 1
        assert body.endswith("= 1")
    elif nm.startswith("y_"):  # one (trivial, r = 1) tile choice per (expert, block, strategy): y[e,n,s,0] = 1
        n_y += 1
        assert re.fullmatch(r"y\[\d+,\d+,\d+,0\] = 1", body)
    elif nm.startswith("x_equal_for"):
        n_eq += 1
        assert body.endswith("= 0")
    else:
        raise SystemExit(f"unexpected constraint {nm}")
d = [[[delta[(e, n, s)] for s in range(S)] for n in range(N)] for e in range(E)]
w = [[[mem[0][(e, n, s)] for s in range(S)] for n in range(N)] for e in range(E)]
json.dump({"source": "bits_model-1.lp (reference repo root), written by mxmoe/quant/bits_solver.py:384",
           "E": E, "N": N, "S": S, "delta": d, "mem_coef": w, "budget": mem[1], "n_one_strategy": n_one, "n_gate_eq_up": n_eq, "n_tile_choice": n_y},
          open(OUT, "w"))
print(E, N, S, mem[1], n_one, n_eq, n_y)
