#!/bin/bash
# same-box A/B of library variants: tools/r02_ab.sh "<variants>" "<workloads>" [rounds]
L=$PWD/mxmoe_b200/lib
export MXB_BENCH_SUSTAINED=0 MXB_BENCH_COPY_ONLY=0
for r in $(seq 1 ${3:-2}); do
 for v in $1; do
  lib=$L/libmxmoe_b200_$v.so; [ "$v" = "default" ] && lib=$L/libmxmoe_b200.so
  for w in $2; do
   MXB_LIB=$lib timeout 300 python bench.py --workload $w --steps 30 --warmup 5 --no-cpu > /tmp/ab.json 2> /tmp/ab.err || { echo "$v $w FAILED"; tail -3 /tmp/ab.err; continue; }
   python - "$v" "$w" <<'PY'
import json, sys
d = json.loads(open("/tmp/ab.json").read().strip().splitlines()[-1]); r = d["roofline"]
print(f"{sys.argv[1]:8s} {sys.argv[2]:4s} step {d['ms_per_step']:.4f} ms  gemm {r.get('ms_per_launch_pair', 0):.4f} ms  frac {r['frac']:.3f}  clk {d['clocks'].get('sm_mhz')}", flush=True)
PY
  done
 done
done
