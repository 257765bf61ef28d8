#!/bin/bash
bash tools/r02_ep.sh 4 c2 c4
MXB_EP_TIMING=0 timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29512 \
   bench.py --gpus 4 --workload c5_stack --steps 5 --warmup 3 > gpurun_out/bench_r02_c5_stack_ep4.json 2> gpurun_out/bench_r02_c5_stack_ep4.err
echo "c5_stack ep4 rc=$?"; tail -1 gpurun_out/bench_r02_c5_stack_ep4.json | cut -c1-400; grep -i "error\|Traceback" -A8 gpurun_out/bench_r02_c5_stack_ep4.err | head -30
