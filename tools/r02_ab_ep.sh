#!/bin/bash
# same-box A/B of library variants on the expert-parallel c2 step: tools/r02_ab_ep.sh N "<variants>" [rounds]
N=$1; L=$PWD/mxmoe_b200/lib
export MXB_BENCH_SUSTAINED=0 MXB_BENCH_COPY_ONLY=0 MXB_EP_TIMING=1
for r in $(seq 1 ${3:-2}); do
 for v in $2; do
  lib=$L/libmxmoe_b200_$v.so; [ "$v" = "default" ] && lib=$L/libmxmoe_b200.so
  MXB_LIB=$lib timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
      bench.py --gpus $N --workload c2 --steps 30 --warmup 5 --no-cpu > /tmp/abep.json 2> /tmp/abep.err || { echo "$v FAILED"; tail -3 /tmp/abep.err; continue; }
  python - "$v" <<'PY'
import json, sys
d = json.loads(open("/tmp/abep.json").read().strip().splitlines()[-1])
print(f"{sys.argv[1]:8s} ep{d['n_gpus']} step {d['ms_per_step']:.4f} ms  {d['value']:.0f} TOPS  parity {d.get('parity_check', {}).get('bit_identical')}", flush=True)
PY
  grep "ep phases rank 0" /tmp/abep.err | cut -c1-200
 done
done
