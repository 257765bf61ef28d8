#!/bin/bash
# usage: tools/bench_ab.sh <workload>...   -> one summary line per workload (device ms, GEMM ms per launch pair, frac)
for w in "$@"; do
  python bench.py --workload $w --steps 30 --warmup 5 --no-cpu > /tmp/b_$w.log 2>&1
  python - "$w" <<'PY'
import sys, json
w = sys.argv[1]
line = [l for l in open(f"/tmp/b_{w}.log") if l.startswith("{")]
if not line:
    print(w, "NO JSON:", open(f"/tmp/b_{w}.log").read()[-400:]); sys.exit(0)
d = json.loads(line[-1]); r = d["roofline"]
print(w, "ms", round(d["ms_per_step"], 4), "e2e_ms", round(d["e2e"]["ms_per_step"], 4), "kernel_ms", r.get("ms_per_launch_pair", r.get("ms_per_launch")), "frac", round(r["frac"], 3))
PY
done
