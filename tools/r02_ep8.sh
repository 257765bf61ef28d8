#!/bin/bash
# the 8-GPU run of round 2: EP parity tests at W=8, c2 / c4 bench lines (+ phase timelines), c2 variants, c5 stack
N=8
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 900 python -m pytest tests/test_gpu_ep.py -q --timeout 600 -k "8-" > gpurun_out/pytest_ep_r02_n8.txt 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_ep_r02_n8.txt
for w in c2 c4; do
  MXB_EP_TIMING=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --workload $w --steps 20 --warmup 5 > gpurun_out/bench_r02_${w}_ep$N.json 2> gpurun_out/bench_r02_${w}_ep$N.err
  echo "bench $w ep$N rc=$?"; tail -1 gpurun_out/bench_r02_${w}_ep$N.json | cut -c1-200; grep "ep phases" gpurun_out/bench_r02_${w}_ep$N.err | head -8
done
run() { tag=$1; shift
  env "$@" MXB_EP_TIMING=1 MXB_BENCH_SUSTAINED=0 MXB_BENCH_COPY_ONLY=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
      bench.py --gpus $N --workload c2 --steps 20 --warmup 5 > gpurun_out/epvar_c2_ep${N}_$tag.json 2> gpurun_out/epvar_c2_ep${N}_$tag.err
  echo "== $tag rc=$?"; tail -1 gpurun_out/epvar_c2_ep${N}_$tag.json | cut -c1-160; grep "ep phases" gpurun_out/epvar_c2_ep${N}_$tag.err | head -2
}
run sms40 MXB_EP_COMM_SMS=40
run bal MXB_EP_PLACEMENT=balanced
run noovl MXB_EP_OVERLAP=0
MXB_BENCH_SUSTAINED=0 timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
   bench.py --gpus $N --workload c5_stack --steps 5 --warmup 3 > gpurun_out/bench_r02_c5_stack_ep8.json 2> gpurun_out/bench_r02_c5_stack_ep8.err
echo "c5_stack ep8 rc=$?"; tail -1 gpurun_out/bench_r02_c5_stack_ep8.json | cut -c1-300; grep -i "error\|Traceback" -A8 gpurun_out/bench_r02_c5_stack_ep8.err | head -20
