#!/usr/bin/env python
"""Per-CTA timeline of w4a16_stream_kernel on the c3 decode step (needs the WS_PROFILE build:
tools/build_variant.sh prof -DWS_PROFILE; MXB_LIB=mxmoe_b200/lib/libmxmoe_b200_prof.so python tools/ws_profile.py [bs])."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import bench
from mxmoe_b200 import config as Cfg, ops
from mxmoe_b200.layer import MoELayer

T = int(sys.argv[1]) if len(sys.argv) > 1 else 64
wl = bench.WORKLOADS["c3"]
s = Cfg.MODEL_SHAPES[wl["model"]]
torch.cuda.set_device(0)
h = ops.Handle(0)
layer = MoELayer(h, s["E"], s["topk"], s["H"], s["I"], s["ns"], bench.qlayer_for(wl["model"], wl["layer"], wl["q"]),
                 max_tokens=T, hadamard_down=False, weight_scale=0.05)
x = (torch.rand(T, s["H"]) * 2 - 1).half().cuda()
ids, w = bench.routing(wl["model"], T, 0, None)
ids, w = ids.cuda(), w.cuda()
out = torch.empty(T, s["H"], dtype=torch.float16, device="cuda")
for _ in range(5):
    layer.forward(x, ids, w, out)
torch.cuda.synchronize()
lib = ops.load_library()
buf = np.zeros((2, 160, 8), dtype=np.uint64)
fn = lib.mxb_debug_ws_profile
fn.restype = C.c_int
fn.argtypes = [C.c_void_p]
assert fn(buf.ctypes.data) == 0
print("rows per expert:", layer.last_counts().tolist() if hasattr(layer.last_counts(), "tolist") else layer.last_counts())
for slot, name in ((0, "gate_up"), (1, "down")):
    p = buf[slot][:148].astype(np.int64)
    t0 = p[:, 0] - p[:, 0].min()
    end = t0 + p[:, 2]
    print(f"== {name}: kernel span {end.max() / 1e3:.1f} us; CTA start spread {t0.max() / 1e3:.1f} us; CTA busy mean {p[:, 2].mean() / 1e3:.1f} "
          f"min {p[:, 2].min() / 1e3:.1f} max {p[:, 2].max() / 1e3:.1f} us; setup mean {p[:, 1].mean() / 1e3:.1f} us; fill mean {p[:, 3].mean() / 1e3:.1f} max {p[:, 3].max() / 1e3:.1f} us")
    stages = p[:, 7] / 3.0  # cost slice in units of 8-token stage costs
    print("   cta  start  busy  fill  ent  two_stages later_stages cost")
    order = np.argsort(-p[:, 2])
    for i in list(order[:12]) + list(order[-8:]):
        print(f"   {i:3d} {t0[i] / 1e3:6.1f} {p[i, 2] / 1e3:6.1f} {p[i, 3] / 1e3:5.1f} {p[i, 4]:4d} {p[i, 5]:6d} {p[i, 6]:6d} {p[i, 7]:7d}")
    # least squares: busy ~ a*ent + b*first_pass_stages + c*later_pass_stages (+ two-tile surcharge)
    tot_stage = np.zeros(148)
    A = np.stack([np.ones(148), p[:, 4], p[:, 5], p[:, 6]], 1).astype(np.float64)
    coef, *_ = np.linalg.lstsq(A, p[:, 2].astype(np.float64), rcond=None)
    print("   lstsq busy_ns ~ const %.0f + %.0f/entry + %.2f/two-tile stage + %.2f/later-pass stage" % tuple(coef))
