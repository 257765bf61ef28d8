"""JSON contracts of the hot path, kept identical to the reference so its files load unchanged.

  QLinearConfig / QExpertConfig / QLayerConfig / QModelConfig, build_qmodel_cfg_from_json
        mxmoe/kernels/qconfig.py:5-80          (schema written by bits_solver.export_qconfig, bits_solver.py:42-53)
  get_qcfg_list                                 run_mxmoe_gg.py:11-29
  load_tile_config (both schemas)               bits_solver.py:30,67-68  and  compose_kernel.py:88-91
  MODEL_SHAPES / MODEL_ID_TO_LAYERS             gen_workload.py:15-20, bits_solver.py:636-644, HF configs
"""
from __future__ import annotations

import dataclasses
import json
import re
from typing import Dict, List, Optional, Tuple


BITS_E4M3 = -8  # MXB_BITS_E4M3 (include/mxmoe_b200.h): 8-bit E4M3 containers


@dataclasses.dataclass
class QLinearConfig:
    w_bits: int = 16
    w_gsize: int = -1
    w_sym: bool = False
    w_clip: Tuple[float, float] = (1.0, 1.0)
    a_bits: int = 16
    a_gsize: int = -1
    a_sym: bool = True
    a_clip: Tuple[float, float] = (1.0, 1.0)

    def __str__(self):
        return self.qstr().upper() if self.w_bits == BITS_E4M3 else f"W{self.w_bits}A{self.a_bits}_g{self.w_gsize}_{'sym' if self.w_sym else 'asym'}"

    def qstr(self) -> str:
        if self.w_bits == 16 and self.a_bits == 16:
            return "fp16"
        if self.w_bits == BITS_E4M3:  # the reference's name for the FP8 kind (tile_config.py:45)
            return f"w8a8_g{self.w_gsize}_sym_E4M3"
        return f"w{self.w_bits}a{self.a_bits}_g{self.w_gsize}_{'sym' if self.w_sym else 'asym'}"

    def to_dict(self) -> dict:
        return dataclasses.asdict(self)

    @staticmethod
    def from_dict(d: dict) -> "QLinearConfig":
        return QLinearConfig(w_bits=d["w_bits"], w_gsize=d["w_gsize"], w_sym=d["w_sym"], w_clip=tuple(d["w_clip"]),
                             a_bits=d["a_bits"], a_gsize=d["a_gsize"], a_sym=d["a_sym"], a_clip=tuple(d["a_clip"]))

    @staticmethod
    def from_qstr(qstr: Optional[str]) -> "QLinearConfig":
        """`w4a16_g128_asym` -> config; None / 'fp16' -> 16/16 (parsing of gen_workload.py:48-60)."""
        if qstr is None or qstr == "fp16":
            return QLinearConfig(16, -1, True, (1.0, 1.0), 16, -1, True, (1.0, 1.0))
        m = re.fullmatch(r"w(\d+)a(\d+)_g(-?\d+)_(sym|asym)(_E4M3)?", qstr)
        if not m:
            raise ValueError(f"bad qstr {qstr!r}")
        w, a, g = int(m.group(1)), int(m.group(2)), int(m.group(3))
        if m.group(5):  # 8-bit float containers: bit-width code MXB_BITS_E4M3 for both operands
            if (w, a, g, m.group(4)) != (8, 8, -1, "sym"):
                raise ValueError(f"bad qstr {qstr!r}: the E4M3 kind is w8a8_g-1_sym_E4M3")
            w = a = BITS_E4M3
        return QLinearConfig(w, g, m.group(4) == "sym", (1.0, 1.0), a, g, True, (1.0, 1.0))


@dataclasses.dataclass
class QExpertConfig:
    gate: QLinearConfig
    up: QLinearConfig
    down: QLinearConfig

    def to_dict(self) -> dict:
        return dataclasses.asdict(self)

    def qmap(self) -> Dict[str, QLinearConfig]:
        return {"gate": self.gate, "up": self.up, "down": self.down}


@dataclasses.dataclass
class QLayerConfig:
    experts: Dict[str, QExpertConfig]

    def to_dict(self) -> dict:
        return dataclasses.asdict(self)


@dataclasses.dataclass
class QModelConfig:
    layers: Dict[str, QLayerConfig]
    LT: Optional[dict] = None

    def to_dict(self) -> dict:
        return {k: v.to_dict() for k, v in self.layers.items()}


def qmodel_cfg_from_dict(d: dict) -> QModelConfig:
    d = dict(d)
    LT = d.pop("LT", None)
    layers = {}
    for layer_idx, layer_cfg in d.items():
        experts = {}
        for expert_idx, ec in layer_cfg["experts"].items():
            experts[expert_idx] = QExpertConfig(gate=QLinearConfig.from_dict(ec["gate"]), up=QLinearConfig.from_dict(ec["up"]),
                                                down=QLinearConfig.from_dict(ec["down"]))
        layers[layer_idx] = QLayerConfig(experts=experts)
    return QModelConfig(layers=layers, LT=LT)


def build_qmodel_cfg_from_json(json_path: str) -> QModelConfig:
    with open(json_path, "r") as f:
        return qmodel_cfg_from_dict(json.load(f))


def build_uni_qexpert_cfg(qcfg: QLinearConfig) -> QExpertConfig:
    return QExpertConfig(gate=qcfg, up=qcfg, down=qcfg)


def build_uni_qlayer_cfg(qcfg: QLinearConfig, num_experts: int) -> QLayerConfig:
    return QLayerConfig(experts={str(i): build_uni_qexpert_cfg(qcfg) for i in range(num_experts)})


def build_uni_qmodel_cfg(qcfg: QLinearConfig, num_layers: int, num_experts: int) -> QModelConfig:
    return QModelConfig(layers={str(i): build_uni_qlayer_cfg(qcfg, num_experts) for i in range(num_layers)})


def get_qcfg_list(qcfg_file: str, target_layer: int) -> set:
    """The set of `w{W}a{A}_g{G}_{sym}` strings a qconfig uses (run_mxmoe_gg.py:11-29)."""
    with open(qcfg_file, "r") as f:
        qcfg = json.load(f)
    qcfg.pop("LT", None)
    out = set()
    for layer_idx, v in qcfg.items():
        if target_layer != -1 and int(layer_idx) != target_layer:
            continue
        for _, qexp in v["experts"].items():
            for _, ql in qexp.items():
                out.add(f"w{ql['w_bits']}a{ql['a_bits']}_g{ql['w_gsize']}_{'sym' if ql['w_sym'] else 'asym'}")
    return out


_TILE_RE = re.compile(r"TileConfig\(([^)]*)\)")


def _parse_tile_repr(s: str) -> List[dict]:
    out = []
    for m in _TILE_RE.finditer(s):
        kv = {}
        for part in m.group(1).split(","):
            if "=" in part:
                k, v = part.split("=", 1)
                try:
                    kv[k.strip()] = int(v.strip())
                except ValueError:
                    kv[k.strip()] = v.strip().strip("'\"")
        out.append(kv)
    return out


def load_tile_config(path: str) -> Dict[str, List[dict]]:
    """Accept both tile_config schemas of the reference:
    {layer_idx: repr(tuple[TileConfig,...])} written by bits_solver.export_qconfig, and
    {qcfg_str: [TileConfig fields...]} read by compose_kernel.get_tile_configs.  Returns {key: [dict,...]}.
    On B200 only BN (tile N of the UMMA) is a free parameter; the other fields are carried for the record."""
    with open(path) as f:
        raw = json.load(f)
    out = {}
    for k, v in raw.items():
        if isinstance(v, str):
            out[k] = _parse_tile_repr(v)
        elif isinstance(v, list):
            out[k] = [x if isinstance(x, dict) else (_parse_tile_repr(x)[0] if isinstance(x, str) and _parse_tile_repr(x) else {"repr": x})
                      for x in v]
        else:
            out[k] = [v]
    return out


def tile_for(tiles: Dict[str, List[dict]], qstr: str, layer: int, qcfg_list: List[str]) -> Optional[dict]:
    """The tile_config entry that applies to problems of kind `qstr` in `layer`, for both schemas:
    {qcfg: [TileConfig...]} (compose_kernel.py:87-91: the first entry is the one the generator instantiates) and
    {layer: (TileConfig per strategy)} (bits_solver.py:30,67-68: by this library's export matched on its QCFG field,
    by the reference's export positionally in the sorted strategy list)."""
    if not tiles:
        return None
    if qstr in tiles:
        return tiles[qstr][0] if tiles[qstr] else None
    ent = tiles.get(str(layer))
    if not ent:
        return None
    for t in ent:
        if t.get("QCFG") == qstr:
            return t
    order = sorted(qcfg_list)
    if qstr in order and order.index(qstr) < len(ent) and "QCFG" not in ent[order.index(qstr)]:
        return ent[order.index(qstr)]
    return None


def tile_hint(tile: Optional[dict]) -> int:
    """MxbProblem.tile_n for a tile_config entry.  On sm_100 the MMA atom is fixed by the hardware, so an entry selects the
    kernel path and the UMMA N: entries written by this library carry PATH (2 = CTA-pair kernel, 1 = streaming kernel:
    both choose the token extent per tile themselves -> 0; 0 = single-CTA kernel at UMMA N = BN); an entry of the
    reference (sm_80/89 tiles, no PATH) is honoured at the granularity that exists here: its BN becomes the UMMA N of the
    single-CTA kernel, rounded down to 32/64/128/256."""
    if not tile or "BN" not in tile:
        return 0
    if tile.get("PATH", 0) in (1, 2):
        return 0
    bn = int(tile["BN"])
    return 256 if bn >= 256 else 128 if bn >= 128 else 64 if bn >= 64 else 32


MODEL_ID_TO_LAYERS = {"qwen2_moe": 24, "qwen2_moe_57b": 28, "ds2": 27, "mixtral": 32}

# E routed experts, top-k, shared-expert multiple ns, hidden H, moe intermediate I (SURVEY §8 shape table)
MODEL_SHAPES = {
    "qwen2_moe": dict(E=60, topk=4, ns=4, H=2048, I=1408, renorm=False),
    "ds2": dict(E=64, topk=6, ns=2, H=2048, I=1408, renorm=False),
    "mixtral": dict(E=8, topk=2, ns=0, H=4096, I=14336, renorm=True),
    "qwen2_moe_57b": dict(E=64, topk=8, ns=8, H=3584, I=2560, renorm=False),
}
