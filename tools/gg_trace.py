#!/usr/bin/env python
"""Per-k-block timeline of gg_pair_kernel's roles on the first CTA pair (needs a -DGP_TRACE build: tools/build_variant.sh trace
-DGP_TRACE, run with MXB_LIB=.../libmxmoe_b200_trace.so).  usage: gg_trace.py [workload] [gu|dn]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from mxmoe_b200 import config as CFG  # noqa: E402
from mxmoe_b200 import ops  # noqa: E402
from mxmoe_b200.layer import MoELayer  # noqa: E402

wname = sys.argv[1] if len(sys.argv) > 1 else "c2"
which = sys.argv[2] if len(sys.argv) > 2 else "gu"
sm_limit = int(sys.argv[3]) if len(sys.argv) > 3 else 0
wl = bench.WORKLOADS[wname]
s = CFG.MODEL_SHAPES[wl["model"]]
T = wl["bs"]
h = ops.Handle(0)
ql = bench.qlayer_for(wl["model"], wl.get("layer", 0), wl["q"])
layer = MoELayer(h, s["E"], s["topk"], s["H"], s["I"], s["ns"], ql, max_tokens=T, hadamard_down=wl["hadamard"], weight_scale=0.05)
ids, w = (t.cuda() for t in bench.routing(wl["model"], T, seed=0, device="cuda"))
x = (torch.rand(T, s["H"], device="cuda") * 2 - 1).half()
out = torch.empty(T, s["H"], dtype=torch.float16, device="cuda")
for _ in range(3):
    layer.forward(x, ids, w, out)
torch.cuda.synchronize()
lib = ops.load_library()
if not hasattr(lib, "mxb_debug_gp_trace"):
    print("forward ok; this library has no trace build (MXB_LIB=.../libmxmoe_b200_trace.so)")
    sys.exit(0)
lib.mxb_debug_gp_trace.argtypes = [C.c_void_p, C.c_int]
N = 2048
buf = np.zeros((9, N), dtype=np.uint64)
lib.mxb_debug_gp_trace(None, 1)
if sm_limit:
    h.set_sm_limit(sm_limit)
    print("SM limit", sm_limit)
h.group_gemm(layer._gu_arr if which == "gu" else layer._dn_arr)
torch.cuda.synchronize()
assert lib.mxb_debug_gp_trace(buf.ctypes.data, 0) == 0
t0 = buf[buf > 0].min()
tr = [(r[r > 0].astype(np.int64) - int(t0)) / 1000.0 for r in buf]  # us
names = ["producer: stage free", "mma: full seen", "mma: conv seen", "mma: issued", "conv: slab landed", "conv: arrived", "epi: tmem_full seen",
         "epi: tile stored", "mma: acc free seen"]
for i, n in enumerate(names):
    print(f"{n:22s} n={len(tr[i]):5d} last {tr[i][-1] if len(tr[i]) else 0:8.1f} us")
nk = min(len(tr[1]), len(tr[3]))
full, conv, iss = tr[1][:nk], tr[2][:nk], tr[3][:nk]
per = np.diff(iss)
print(f"k-blocks {nk}: mean period {per.mean()*1000:.0f} ns, median {np.median(per)*1000:.0f} ns, p10 {np.percentile(per,10)*1000:.0f}, p90 {np.percentile(per,90)*1000:.0f}")
print(f"  wait full  (prev issued -> full seen): mean {np.mean(full[1:] - iss[:-1])*1000:.0f} ns, median {np.median(full[1:] - iss[:-1])*1000:.0f}")
print(f"  wait conv  (full seen -> conv seen):   mean {np.mean(conv - full)*1000:.0f} ns, median {np.median(conv - full)*1000:.0f}")
print(f"  issue      (conv seen -> issued):      mean {np.mean(iss - conv)*1000:.0f} ns, median {np.median(iss - conv)*1000:.0f}")
nc = min(len(tr[4]), len(tr[5]))
if nc:
    print(f"converter: landed -> arrived mean {np.mean(tr[5][:nc] - tr[4][:nc])*1000:.0f} ns; period mean {np.diff(tr[5][:nc]).mean()*1000:.0f} ns")
ne = min(len(tr[6]), len(tr[7]))
if ne:
    print(f"epilogue: tiles {ne}, drain (tmem_full seen -> stored) mean {np.mean(tr[7][:ne] - tr[6][:ne]):.2f} us; tile period {np.diff(tr[7][:ne]).mean():.2f} us")
print("first 40 k-blocks: (full, conv, issued) us")
for i in list(range(min(24, nk))) + list(range(200, min(224, nk))):
    print(f"  kb {i:3d}  prod_free {tr[0][i] if i < len(tr[0]) else -1:7.2f}  slab_landed {tr[4][i] if i < len(tr[4]) else -1:7.2f}  conv_arrived {tr[5][i] if i < len(tr[5]) else -1:7.2f}"
          f"  full {full[i]:7.2f}  conv {conv[i]:7.2f}  issued {iss[i]:7.2f}  stage_free_again {tr[0][i + 6] if i + 6 < len(tr[0]) else -1:7.2f}")
print("tiles: acc free seen / tmem_full seen / stored")
for i in range(min(12, ne)):
    print(f"  tile {i:2d}  acc_free {tr[8][i] if i < len(tr[8]) else -1:7.2f}  tmem_full {tr[6][i]:7.2f}  stored {tr[7][i]:7.2f}")
