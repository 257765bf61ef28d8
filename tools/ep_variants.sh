#!/bin/bash
# tools/ep_variants.sh <ngpus> <workload>: the EP bench under a few overlap settings (run under gpurun --gpus N)
N=$1; WL=${2:-c2}
run() {
  tag=$1; shift
  env "$@" MXB_EP_TIMING=1 timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --workload $WL --steps 20 --warmup 5 > gpurun_out/ep_${tag}.json 2> gpurun_out/ep_${tag}.err
  python - <<PY
import json
for l in open("gpurun_out/ep_${tag}.json"):
    if l.startswith("{"):
        d = json.loads(l); print("${tag}", round(d["ms_per_step"], 4), round(d["value"], 1), "e2e", round(d["e2e"]["ms_per_step"], 4))
PY
  grep "ep phases" gpurun_out/ep_${tag}.err; grep -m2 "Error" gpurun_out/ep_${tag}.err | cut -c1-200
}
run tail1 MXB_EP_TAIL_WAIT=1
run tail0 MXB_EP_TAIL_WAIT=0
run noovl MXB_EP_OVERLAP=0
run tail0_sm32 MXB_EP_TAIL_WAIT=0 MXB_EP_COMM_SMS=32
