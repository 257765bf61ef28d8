"""On-disk format for quantized experts in the library's native layouts, with loader and converter.

The reference keeps quantized models as `torch.save` of FAKE-quantized fp16 state dicts (mxmoe/quant/quant.py:772-794):
16 bits per weight on disk and a full re-quantization + re-pack (quantize.cuh:218-475) at every start.  Here the packed
int4/int8/E4M3 bytes and their fp16 scales — exactly the operands `MxbProblem.B` / `scale_b` point at — are stored once,
next to the `QModelConfig` that produced them, and a layer starts by copying them to HBM.

File layout (little endian), one file per model or per layer:

    bytes 0..7    magic  b"MXBPACK1"
    bytes 8..15   u64    length L of the JSON header
    bytes 16..    L bytes of UTF-8 JSON, zero-padded so that the data section starts at a multiple of 64
    data section  tensors back to back, each starting at a multiple of 64 (offsets in the header are relative to it)

    header = {"format": 1, "meta": {...free form...},
              "layers": {"<layer index>": {"E": .., "top_k": .., "H": .., "I": .., "ns": .., "hadamard_down": bool,
                                           "gate_up_interleaved": bool,
                                           "experts": {"<e>": {"gate_up": "<qstr>", "down": "<qstr>",
                                                               "tensors": {"gu_packed": T, "gu_scale": T|null,
                                                                           "dn_packed": T, "dn_scale": T|null}}}}}}
    T = {"dtype": "uint8"|"float16"|.., "shape": [...], "offset": int, "nbytes": int, "crc32": int}

Expert index E is the fused shared expert, as everywhere in this package.  `gate_up_interleaved` records whether gate/up
rows are stored pairwise adjacent (the layout the fused silu*mul epilogue needs, MoELayer.fused_silu); the loader refuses
a file whose interleaving does not match the layer it is loaded into rather than silently computing something else.
"""
from __future__ import annotations

import json
import os
import struct
import zlib
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from .config import QLayerConfig, QLinearConfig, QExpertConfig

MAGIC = b"MXBPACK1"
ALIGN = 64
_TENSORS = ("gu_packed", "gu_scale", "dn_packed", "dn_scale")
_DTYPES = {"uint8": torch.uint8, "int8": torch.int8, "float16": torch.float16, "int32": torch.int32, "float32": torch.float32}
_DTYPE_NAMES = {v: k for k, v in _DTYPES.items()}


class CheckpointError(ValueError):
    pass


def _pad(n: int) -> int:
    return (n + ALIGN - 1) // ALIGN * ALIGN


def _bytes_of(t: torch.Tensor) -> bytes:
    t = t.detach().contiguous().cpu()
    return t.view(torch.uint8).numpy().tobytes() if t.numel() else b""


def layer_record(layer) -> dict:
    """The serialisable description of a MoELayer (mxmoe_b200/layer.py) and its packed experts."""
    ids = layer.local_experts + ([layer.E] if layer.ns else [])
    experts = {}
    for e, ew in zip(ids, layer.experts):
        experts[str(e)] = {"gate_up": ew.q_gu.qstr(), "down": ew.q_dn.qstr(),
                           "tensors": {"gu_packed": ew.gu_packed, "gu_scale": ew.gu_scale, "dn_packed": ew.dn_packed,
                                       "dn_scale": ew.dn_scale}}
    return {"E": layer.E, "top_k": layer.topk, "H": layer.H, "I": layer.I, "ns": layer.ns, "hadamard_down": bool(layer.hadamard_down),
            "gate_up_interleaved": bool(layer.fused_silu or getattr(layer, "stream_epi", False)), "experts": experts}


def save_packed(path: str, layers: Dict[int, dict], meta: Optional[dict] = None) -> int:
    """layers: {layer index: layer_record(...)-shaped dict whose "tensors" values are torch tensors (or None)}.
    Returns the number of bytes written."""
    blobs, off = [], 0
    header = {"format": 1, "meta": meta or {}, "layers": {}}
    for li, rec in layers.items():
        out = {k: v for k, v in rec.items() if k != "experts"}
        out["experts"] = {}
        for e, er in rec["experts"].items():
            tens = {}
            for name in _TENSORS:
                t = er["tensors"].get(name)
                if t is None:
                    tens[name] = None
                    continue
                if t.dtype not in _DTYPE_NAMES:
                    raise CheckpointError(f"layer {li} expert {e} {name}: dtype {t.dtype} is not storable")
                b = _bytes_of(t)
                tens[name] = {"dtype": _DTYPE_NAMES[t.dtype], "shape": list(t.shape), "offset": off, "nbytes": len(b),
                              "crc32": zlib.crc32(b) & 0xFFFFFFFF}
                blobs.append((off, b))
                off = _pad(off + len(b))
            out["experts"][str(e)] = {"gate_up": er["gate_up"], "down": er["down"], "tensors": tens}
        header["layers"][str(li)] = out
    hj = json.dumps(header, sort_keys=True).encode("utf-8")
    data_start = _pad(16 + len(hj))
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(MAGIC)
        f.write(struct.pack("<Q", len(hj)))
        f.write(hj)
        f.write(b"\0" * (data_start - 16 - len(hj)))
        pos = 0
        for o, b in blobs:
            if o > pos:
                f.write(b"\0" * (o - pos))
            f.write(b)
            pos = o + len(b)
        total = data_start + pos
    os.replace(tmp, path)  # a crashed writer never leaves a half-written file under the final name
    return total


def read_header(path: str) -> Tuple[dict, int]:
    """Returns (header, offset of the data section)."""
    with open(path, "rb") as f:
        head = f.read(16)
        if len(head) < 16 or head[:8] != MAGIC:
            raise CheckpointError(f"{path}: not a packed-expert file (bad magic)")
        (n,) = struct.unpack("<Q", head[8:])
        size = os.fstat(f.fileno()).st_size
        if 16 + n > size:
            raise CheckpointError(f"{path}: truncated header ({n} bytes announced, file has {size})")
        try:
            header = json.loads(f.read(n).decode("utf-8"))
        except (UnicodeDecodeError, json.JSONDecodeError) as ex:
            raise CheckpointError(f"{path}: unreadable header: {ex}") from None
    if header.get("format") != 1:
        raise CheckpointError(f"{path}: unknown format version {header.get('format')!r}")
    return header, _pad(16 + n)


def load_packed(path: str, device="cpu", layers=None, verify: bool = True) -> Dict[int, dict]:
    """Reads the file back into {layer index: record} with torch tensors on `device`.  The file is memory-mapped and each
    tensor goes host -> device in one copy (no intermediate fp16 weights, no re-quantization).  layers: optional subset."""
    header, base = read_header(path)
    size = os.path.getsize(path)
    mm = np.memmap(path, dtype=np.uint8, mode="r") if size > 0 else None
    out = {}
    for li, rec in header["layers"].items():
        if layers is not None and int(li) not in layers:
            continue
        r = {k: v for k, v in rec.items() if k != "experts"}
        r["experts"] = {}
        for e, er in rec["experts"].items():
            tens = {}
            for name in _TENSORS:
                d = er["tensors"].get(name)
                if d is None:
                    tens[name] = None
                    continue
                lo, hi = base + d["offset"], base + d["offset"] + d["nbytes"]
                if hi > size:
                    raise CheckpointError(f"{path}: layer {li} expert {e} {name} runs past the end of the file")
                raw = np.array(mm[lo:hi])  # one host copy out of the page cache
                if verify and (zlib.crc32(raw.tobytes()) & 0xFFFFFFFF) != d["crc32"]:
                    raise CheckpointError(f"{path}: layer {li} expert {e} {name}: checksum mismatch")
                dt = _DTYPES[d["dtype"]]
                t = torch.from_numpy(raw).view(dt).reshape(d["shape"]) if d["nbytes"] else torch.empty(d["shape"], dtype=dt)
                tens[name] = t.to(device)
            r["experts"][e] = {"gate_up": er["gate_up"], "down": er["down"], "tensors": tens}
        out[int(li)] = r
    return out


def qlayer_of(record: dict) -> QLayerConfig:
    """The per-expert qconfig of a stored layer, in the reference's schema (gate == up: one fused GEMM)."""
    ex = {}
    for e, er in record["experts"].items():
        gu, dn = QLinearConfig.from_qstr(er["gate_up"]), QLinearConfig.from_qstr(er["down"])
        ex[str(e)] = QExpertConfig(gate=gu, up=gu, down=dn)
    return QLayerConfig(experts=ex)


def layer_from_record(handle, record: dict, max_tokens: int, expert_ids=None, device=None):
    """Builds a MoELayer straight from stored operands (no fp16 weights, no quantization pass)."""
    from .layer import MoELayer
    return MoELayer(handle, record["E"], record["top_k"], record["H"], record["I"], record["ns"], qlayer_of(record), max_tokens=max_tokens,
                    hadamard_down=record["hadamard_down"], expert_ids=expert_ids, device=device, packed=record)


def convert_fake_quant_state_dict(handle, state_dict: Dict[str, torch.Tensor], qlayer: QLayerConfig, layer_idx: int, E: int, top_k: int,
                                  max_tokens: int = 1, hadamard_down: bool = False,
                                  expert_fmt: str = "model.layers.{L}.mlp.experts.{e}.{proj}.weight",
                                  shared_fmt: Optional[str] = "model.layers.{L}.mlp.shared_expert.{proj}.weight") -> dict:
    """One decoder layer of a reference checkpoint -> a storable record.

    state_dict holds the fake-quantized fp16 `gate_proj` / `up_proj` / `down_proj` weights the reference saves
    (quant.py:772-794; GPTQ(-had) results from gptq.py:179-275 included — they are already on their grids).  Each expert
    is re-quantized with its own qconfig and packed by the device kernel (`mxb_quant_pack_weight`): a weight that lies on
    a (scale, zero) grid is mapped back to its code, up to the one-ulp rounding of the recomputed fp16 scale."""
    from .layer import MoELayer
    weights = {}
    H = I = None
    for e in range(E):
        g, u, d = (state_dict[expert_fmt.format(L=layer_idx, e=e, proj=p)] for p in ("gate_proj", "up_proj", "down_proj"))
        H, I = d.shape[0], g.shape[0]
        weights[e] = (torch.cat([g, u], 0).half(), d.half())
    ns = 0
    if shared_fmt is not None and shared_fmt.format(L=layer_idx, proj="gate_proj") in state_dict:
        g, u, d = (state_dict[shared_fmt.format(L=layer_idx, proj=p)] for p in ("gate_proj", "up_proj", "down_proj"))
        if g.shape[0] % I:
            raise CheckpointError("the shared expert's intermediate size must be a multiple of the routed one")
        ns = g.shape[0] // I
        weights[E] = (torch.cat([g, u], 0).half(), d.half())
    layer = MoELayer(handle, E, top_k, H, I, ns, qlayer, max_tokens=max_tokens, hadamard_down=hadamard_down, weights=weights)
    return layer_record(layer)
