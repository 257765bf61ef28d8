#!/bin/bash
# EP variants at N GPUs: bash tools/r02_epvar.sh N workload
N=$1; w=$2
mkdir -p gpurun_out
run() { tag=$1; shift
  env "$@" MXB_EP_TIMING=1 MXB_BENCH_SUSTAINED=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
      bench.py --gpus $N --workload $w --steps 20 --warmup 5 > gpurun_out/epvar_${w}_ep${N}_$tag.json 2> gpurun_out/epvar_${w}_ep${N}_$tag.err
  echo "== $tag rc=$?"; python - <<PY
import json
d=json.loads(open("gpurun_out/epvar_${w}_ep${N}_$tag.json").read().strip().splitlines()[-1])
print(round(d["ms_per_step"],4), "ms", round(d["value"],1), "TOPS e2e", round(d["e2e"]["ms_per_step"],3), d.get("parity_check",{}).get("bit_identical"))
PY
  grep "ep phases" gpurun_out/epvar_${w}_ep${N}_$tag.err | head -2
}
run base MXB_X=0
run noovl MXB_EP_OVERLAP=0
run noovl_bal MXB_EP_OVERLAP=0 MXB_EP_PLACEMENT=balanced
run bal MXB_EP_PLACEMENT=balanced
