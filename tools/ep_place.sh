#!/bin/bash
# tools/ep_place.sh <ngpus>: EP bench (c2) with popularity-balanced and contiguous expert placement
N=$1
for pl in balanced contiguous; do
  MXB_EP_PLACEMENT=$pl MXB_EP_TIMING=1 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/ep${N}_${pl}.json 2> gpurun_out/ep${N}_${pl}.err
  python - <<PY
import json
for l in open("gpurun_out/ep${N}_${pl}.json"):
    if l.startswith("{"):
        d = json.loads(l); print("${N} ${pl}", round(d["ms_per_step"], 4), round(d["value"], 1), "e2e", round(d["e2e"]["ms_per_step"], 4))
PY
  grep "ep phases" gpurun_out/ep${N}_${pl}.err; grep -m2 "Error" gpurun_out/ep${N}_${pl}.err | cut -c1-200
done
