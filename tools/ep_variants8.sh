#!/bin/bash
# tools/ep_variants8.sh <ngpus> <workload>: default (overlap) and no-overlap EP bench
N=$1; WL=${2:-c2}
run() {
  tag=$1; shift
  env "$@" MXB_EP_TIMING=1 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --workload $WL --steps 20 --warmup 5 > gpurun_out/ep${N}_${WL}_${tag}.json 2> gpurun_out/ep${N}_${WL}_${tag}.err
  python - <<PY
import json
for l in open("gpurun_out/ep${N}_${WL}_${tag}.json"):
    if l.startswith("{"):
        d = json.loads(l); print("${N} ${WL} ${tag}", round(d["ms_per_step"], 4), round(d["value"], 1), "e2e", round(d["e2e"]["ms_per_step"], 4))
PY
  grep "ep phases" gpurun_out/ep${N}_${WL}_${tag}.err; grep -m2 "Error" gpurun_out/ep${N}_${WL}_${tag}.err | cut -c1-200
}
run ovl MXB_EP_OVERLAP=1
run noovl MXB_EP_OVERLAP=0
