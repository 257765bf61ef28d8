"""Packed-expert file format (mxmoe_b200/checkpoint.py): CPU-side round trips and error behaviour."""
import json
import struct

import numpy as np
import pytest
import torch

from mxmoe_b200 import checkpoint as CK


def _record(seed=0, E=3, H=64, I=32, ns=1, interleaved=False):
    rng = np.random.default_rng(seed)

    def u8(*s):
        return torch.from_numpy(rng.integers(0, 256, s, dtype=np.uint8))

    def h(*s):
        return torch.from_numpy(rng.random(s).astype(np.float16))

    experts = {}
    for e in range(E + (1 if ns else 0)):
        Ie = I * ns if e == E else I
        if e % 3 == 0:    # w4a4 sym per-channel: 4-bit packed, one scale per row
            gu, gs, dn, ds, q = u8(2 * Ie, H // 2), h(2 * Ie), u8(H, Ie // 2), h(H), "w4a4_g-1_sym"
        elif e % 3 == 1:  # w8a8
            gu, gs, dn, ds, q = u8(2 * Ie, H).view(torch.int8), h(2 * Ie), u8(H, Ie).view(torch.int8), h(H), "w8a8_g-1_sym"
        else:             # fp16: no scales
            gu, gs, dn, ds, q = h(2 * Ie, H), None, h(H, Ie), None, "fp16"
        experts[str(e)] = {"gate_up": q, "down": q, "tensors": {"gu_packed": gu, "gu_scale": gs, "dn_packed": dn, "dn_scale": ds}}
    return {"E": E, "top_k": 2, "H": H, "I": I, "ns": ns, "hadamard_down": False, "gate_up_interleaved": interleaved, "experts": experts}


def _same(a, b):
    assert {k: v for k, v in a.items() if k != "experts"} == {k: v for k, v in b.items() if k != "experts"}
    assert a["experts"].keys() == b["experts"].keys()
    for e in a["experts"]:
        ea, eb = a["experts"][e], b["experts"][e]
        assert (ea["gate_up"], ea["down"]) == (eb["gate_up"], eb["down"])
        for k, ta in ea["tensors"].items():
            tb = eb["tensors"][k]
            if ta is None:
                assert tb is None
            else:
                assert ta.dtype == tb.dtype and ta.shape == tb.shape and torch.equal(ta.view(torch.uint8), tb.view(torch.uint8))


def test_round_trip_two_layers(tmp_path):
    recs = {3: _record(1), 11: _record(2, E=5, ns=0, interleaved=True)}
    p = str(tmp_path / "m.mxb")
    n = CK.save_packed(p, recs, meta={"model": "unit", "wbits": 5.0})
    assert n == (tmp_path / "m.mxb").stat().st_size
    back = CK.load_packed(p)
    assert sorted(back) == [3, 11]
    for li in recs:
        _same(recs[li], back[li])
    only = CK.load_packed(p, layers=[11])
    assert list(only) == [11]
    header, base = CK.read_header(p)
    assert header["meta"] == {"model": "unit", "wbits": 5.0} and base % CK.ALIGN == 0
    for rec in header["layers"].values():  # every tensor starts on a 64-byte boundary of the file
        for er in rec["experts"].values():
            for d in er["tensors"].values():
                assert d is None or (base + d["offset"]) % CK.ALIGN == 0


def test_qlayer_of_matches_stored_strings():
    ql = CK.qlayer_of(_record(0))
    assert ql.experts["0"].gate.qstr() == "w4a4_g-1_sym" and ql.experts["1"].down.qstr() == "w8a8_g-1_sym"
    assert ql.experts["2"].gate.w_bits == 16 and ql.experts["0"].gate is ql.experts["0"].up


def test_empty_tensor_and_no_layers(tmp_path):
    p = str(tmp_path / "e.mxb")
    CK.save_packed(p, {})
    assert CK.load_packed(p) == {}
    rec = _record(0, E=1, ns=0)
    rec["experts"]["0"]["tensors"]["gu_scale"] = torch.empty(0, dtype=torch.float16)
    CK.save_packed(p, {0: rec})
    assert CK.load_packed(p)[0]["experts"]["0"]["tensors"]["gu_scale"].numel() == 0


def test_errors(tmp_path):
    p = str(tmp_path / "m.mxb")
    CK.save_packed(p, {0: _record(0)})
    raw = bytearray(open(p, "rb").read())
    bad = str(tmp_path / "bad.mxb")

    open(bad, "wb").write(b"NOTMAGIC" + bytes(raw[8:]))
    with pytest.raises(CK.CheckpointError, match="bad magic"):
        CK.load_packed(bad)

    open(bad, "wb").write(bytes(raw[:-100]))  # data section cut short
    with pytest.raises(CK.CheckpointError, match="past the end"):
        CK.load_packed(bad)

    flipped = bytearray(raw)
    flipped[-1] ^= 0xFF  # a bit flip inside the last tensor
    open(bad, "wb").write(bytes(flipped))
    with pytest.raises(CK.CheckpointError, match="checksum"):
        CK.load_packed(bad)
    CK.load_packed(bad, verify=False)  # explicit opt-out only

    open(bad, "wb").write(bytes(raw[:8]) + struct.pack("<Q", 1 << 40) + bytes(raw[16:]))
    with pytest.raises(CK.CheckpointError, match="truncated header"):
        CK.load_packed(bad)

    (n,) = struct.unpack("<Q", raw[8:16])
    hdr = json.loads(raw[16:16 + n])
    hdr["format"] = 2
    hj = json.dumps(hdr, sort_keys=True).encode()
    assert len(hj) == n
    open(bad, "wb").write(bytes(raw[:16]) + hj + bytes(raw[16 + n:]))
    with pytest.raises(CK.CheckpointError, match="format version"):
        CK.load_packed(bad)

    rec = _record(0)
    rec["experts"]["0"]["tensors"]["gu_packed"] = torch.zeros(2, dtype=torch.bool)
    with pytest.raises(CK.CheckpointError, match="not storable"):
        CK.save_packed(bad, {0: rec})
