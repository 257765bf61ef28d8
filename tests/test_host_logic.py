"""Host-side logic (JSON contracts, workload generation) against fixtures produced by the reference itself."""
import json
import os

import pytest

from mxmoe_b200 import config as C
from mxmoe_b200 import workload as W

G = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("name,kw", [("fp16", {}), ("qstr", {"qstr": "w4a16_g128_asym"}), ("qcfg", {"qconfig": "file"})])
def test_workload_matches_reference_generator(name, kw):
    """generate_workload_from_gate_trace (gen_workload.py:38-110) run by tests/golden/gen_golden.py"""
    trace = json.load(open(os.path.join(G, "gate_trace.json")))
    gold = json.load(open(os.path.join(G, "workload.json")))[name]
    if "qconfig" in kw:
        kw = {"qconfig": json.load(open(os.path.join(G, "qconfig_mixed.json")))}
    got = W.generate_workload_from_gate_trace(trace, 8192, 11, **kw)
    assert got == gold


def test_qconfig_json_roundtrip_and_qcfg_list():
    path = os.path.join(G, "qconfig_mixed.json")
    qm = C.build_qmodel_cfg_from_json(path)
    raw = json.load(open(path))
    assert set(qm.layers) == {"10", "11"} and qm.LT == raw["LT"]
    e0 = qm.layers["11"].experts["0"]
    assert e0.gate.to_dict() == {**raw["11"]["experts"]["0"]["gate"], "w_clip": (1.0, 1.0), "a_clip": (1.0, 1.0)}
    assert str(e0.gate) == f"W{e0.gate.w_bits}A{e0.gate.a_bits}_g-1_sym"
    assert C.get_qcfg_list(path, 11) == {"w4a4_g-1_sym", "w8a8_g-1_sym"}
    assert C.QLinearConfig.from_qstr("w4a16_g128_asym").w_sym is False
    assert C.QLinearConfig.from_qstr(None).qstr() == "fp16"
    with pytest.raises(ValueError):
        C.QLinearConfig.from_qstr("w4a16")


def test_tile_config_both_schemas(tmp_path):
    a = {"11": "(TileConfig(BM=128, BN=128, BK=64, WM=64, WN=32, WK=64, STAGE=3, SPLITK=1, MMA='MMA_S8_K32'), "
               "TileConfig(BM=256, BN=128, BK=128, WM=64, WN=64, WK=128, STAGE=3, SPLITK=1, MMA='MMA_S4_K64'))"}
    b = {"w8a8_g-1_sym": [{"BM": 128, "BN": 256, "BK": 128, "num_warps": 8}]}
    pa, pb = tmp_path / "a.json", tmp_path / "b.json"
    pa.write_text(json.dumps(a)), pb.write_text(json.dumps(b))
    ta, tb = C.load_tile_config(str(pa)), C.load_tile_config(str(pb))
    assert [t["BN"] for t in ta["11"]] == [128, 128] and ta["11"][1]["MMA"] == "MMA_S4_K64"
    assert tb["w8a8_g-1_sym"][0]["BN"] == 256


def test_synthetic_traces_and_mixed_qconfig():
    tr = W.synthetic_gate_trace("qwen2_moe", [11], "fixture")
    wl = W.generate_workload_from_gate_trace(tr, 8192, 11)
    ms = [p["shape"][0] for p in wl["layer-11"]["gate_up"]]
    assert len(ms) == 61 and ms[-1] == 8192 and 32000 < sum(ms[:-1]) <= 32768
    assert wl["layer-11"]["gate_up"][0]["shape"][1:] == [2816, 2048] and wl["layer-11"]["down"][-1]["shape"] == [8192, 2048, 5632]
    assert W.workload_ops(wl) == sum(2 * p["shape"][0] * p["shape"][1] * p["shape"][2] for part in ("gate_up", "down") for p in wl["layer-11"][part])
    for kind in ("uniform", "zipf"):
        t = W.synthetic_gate_trace("mixtral", [0], kind)
        assert len(t["layer-0"]["access_freq"]) == 8 and t["num_shared_experts"] == 0
    q = W.synthetic_mixed_qconfig("qwen2_moe", [11], 5.0)
    ex = q["11"]["experts"]
    assert len(ex) == 61 and all(e["gate"] == e["up"] for e in ex.values())
    tot = w8 = 0
    for i, e in ex.items():
        sc = 4 if i == "60" else 1
        for blk, wt in (("gate", 2), ("down", 1)):
            tot += wt * sc
            w8 += wt * sc * (e[blk]["w_bits"] == 8)
    assert abs(4 + 4 * w8 / tot - 5.0) < 0.1
    C.qmodel_cfg_from_dict(q)


def test_run_mxmoe_gg_cli_dry_run(tmp_path):
    """Same CLI as the reference driver (run_mxmoe_gg.py:34-41); --dry-run writes only the workload JSON."""
    import subprocess
    import sys
    root = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
    r = subprocess.run([sys.executable, os.path.join(root, "run_mxmoe_gg.py"), "--model", "qwen2_moe", "--dataset", "wiki2",
                        "--bs", "8192", "--layer", "11", "--qconfig", os.path.join(G, "qconfig_mixed.json"), "--dry-run"],
                       capture_output=True, text=True, cwd=str(tmp_path))
    assert r.returncode == 0, r.stderr
    assert "w4a4_g-1_sym" in r.stdout and "w8a8_g-1_sym" in r.stdout
    wl = json.load(open(os.path.join(root, "out", "workloads", "qwen2_moe-wiki2-8192-qcfg.json")))
    assert wl["num_tokens"] == 8192 and len(wl["layer-11"]["gate_up"]) == 61
    assert {p["w_bits"] for p in wl["layer-11"]["down"]} == {4, 8}


def test_gate_up_scale_interleave_roundtrip():
    """Group-major scales ([K/g, N] / [K/g, N, 2], ref quantize.cuh:299-315 permute_scale) follow the gate/up row interleave of
    MXB_EPI_SILU_MUL inside every group row; the inverse restores the stacked order (decode records loaded into large layers)."""
    import torch
    from mxmoe_b200.layer import _interleave_group_major as il
    G, N = 3, 8
    for asym in (False, True):
        c = 2 if asym else 1
        s = torch.arange(G * N * c, dtype=torch.float32).half()
        o = il(s, N, asym)
        v, w = s.view(G, N, c), o.view(G, N, c)
        for i in range(N // 2):
            assert torch.equal(w[:, 2 * i], v[:, i]) and torch.equal(w[:, 2 * i + 1], v[:, N // 2 + i])
        assert torch.equal(il(o, N, asym, inverse=True), s)
    assert il(None, N, False) is None


def test_tile_config_entries_map_to_problem_hints(tmp_path):
    """--tile_config is honoured (config.tile_for / tile_hint): both reference schemas and this library's own export."""
    import json
    from mxmoe_b200 import config as C
    qcfgs = ["w4a4_g-1_sym", "w8a8_g-1_sym"]
    # schema of compose_kernel.get_tile_configs (compose_kernel.py:87-91): {qcfg: [tiles]}, reference (sm_89) tile reprs
    ref_repr = ("TileConfig(BM=128, BN=128, BK=128, WM=2, WN=2, WK=1, STAGE=3, SPLITK=-1, MMA='MMA_S4_K64', "
                "QCFGA=QConfig(QBITS=4, GSIZE=-1, SYM=True), QCFGB=QConfig(QBITS=4, GSIZE=-1, SYM=True))")
    p = tmp_path / "t.json"
    json.dump({"w4a4_g-1_sym": [ref_repr], "w8a8_g-1_sym": [{"BM": 64, "BN": 96, "BK": 64}]}, open(p, "w"))
    t = C.load_tile_config(str(p))
    assert C.tile_hint(C.tile_for(t, "w4a4_g-1_sym", 11, qcfgs)) == 128   # reference tile: BN -> UMMA N of the single-CTA kernel
    assert C.tile_hint(C.tile_for(t, "w8a8_g-1_sym", 11, qcfgs)) == 64    # rounded down to 32/64/128/256
    assert C.tile_for(t, "fp16", 11, qcfgs) is None and C.tile_hint(None) == 0
    # schema of bits_solver.export_qconfig (bits_solver.py:30,67-68): {layer: repr(tuple)}; the reference's is positional in
    # the sorted strategy list, this library's carries QCFG / PATH
    json.dump({"11": "(" + ref_repr + ", " + ref_repr.replace("BN=128", "BN=256") + ")",
               "12": "(TileConfig(BM=256, BN=192, BK=128, WM=1, WN=1, WK=1, STAGE=6, SPLITK=-1, MMA='UMMA_I8_2CTA', PATH=2, "
                     "QCFG='w8a8_g-1_sym'), TileConfig(BM=128, BN=64, BK=128, WM=1, WN=1, WK=1, STAGE=3, SPLITK=-1, MMA='UMMA_I8', "
                     "PATH=0, QCFG='w4a4_g128_sym'))"}, open(p, "w"))
    t = C.load_tile_config(str(p))
    assert C.tile_hint(C.tile_for(t, "w4a4_g-1_sym", 11, qcfgs)) == 128
    assert C.tile_hint(C.tile_for(t, "w8a8_g-1_sym", 11, qcfgs)) == 256
    assert C.tile_hint(C.tile_for(t, "w8a8_g-1_sym", 12, qcfgs)) == 0     # PATH 2: the CTA-pair kernel picks its token tile
    assert C.tile_hint(C.tile_for(t, "w4a4_g128_sym", 12, qcfgs)) == 64
    assert C.tile_for(t, "w4a4_g-1_sym", 12, qcfgs) is None


def test_solver_exports_tile_cfg_next_to_qconfig(tmp_path):
    """bits_solver writes `<out>_tile_cfg.json` like the reference's export_qconfig (bits_solver.py:67-68), from mxb_select_tiles."""
    import json
    from mxmoe_b200 import bits_solver as B
    from mxmoe_b200 import config as C
    out = tmp_path / "q.json"
    B.main(["--model", "qwen2_moe", "--wbits", "5.0", "--layers", "3", "--out", str(out)])
    t = C.load_tile_config(str(tmp_path / "q_tile_cfg.json"))
    assert set(t) == {"3"} and {e["QCFG"] for e in t["3"]} == {"w4a4_g-1_sym", "w8a8_g-1_sym"}
    for e in t["3"]:
        assert e["PATH"] == 2 and e["BM"] == 256 and e["BN"] % 16 == 0 and e["MMA"] == "UMMA_I8_2CTA"
