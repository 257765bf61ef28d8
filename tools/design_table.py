#!/usr/bin/env python
"""Print DESIGN.md section 3.3's table rows from the bench lines under profiles/ (bench_<round>_<cfg>.json; round = argv[1], default r02)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RND = sys.argv[1] if len(sys.argv) > 1 else "r02"
NAMES = {"c1": "c1 Qwen1.5 L11 fp16 bs=512", "c2": "**c2** Qwen1.5 L11 w4a4+w8a8 5.0 bit +Hadamard bs=8192",
         "c3": "c3 DeepSeek-V2-Lite w4a16_g128 bs=64", "c4": "c4 Mixtral w8a8 bs=8192",
         "c4_e4m3": "c4 shape, w8a8 E4M3 (kind::f8f6f4)", "c5": "c5 Qwen2-57B w4a4+w8a8 +Hadamard bs=8192"}
for c, name in NAMES.items():
    p = os.path.join(ROOT, "profiles", f"bench_{RND}_{c}.json")
    if not os.path.exists(p):
        continue
    lines = [l for l in open(p) if l.startswith("{")]
    if not lines:
        continue
    d = json.loads(lines[-1])
    r, cb = d["roofline"], d.get("cpu_baseline") or {}
    kern = r["kernel"].split(" (")[0]
    print(f"| {name} | {d['ms_per_step']:.3f} | {d['value']:.0f} | {d['e2e']['value']:.0f} | `{kern}`, {r['frac']:.2f} of "
          f"{'HBM copy' if r['bound'] == 'hbm' else '2x bf16'} | {cb.get('value', float('nan')):.2f} |")
