#!/bin/bash
# round 2 expert-parallel run on N GPUs of one box:  bash tools/r02_ep.sh N [workloads...]
N=$1; shift; WL=${@:-c2 c4}
mkdir -p gpurun_out
nvidia-smi -L | head -8
timeout 1500 python -m pytest tests/test_gpu_ep.py -q --timeout 600 > gpurun_out/pytest_ep_r02_n$N.txt 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_ep_r02_n$N.txt
for w in $WL; do
  MXB_EP_TIMING=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --workload $w --steps 20 --warmup 5 > gpurun_out/bench_r02_${w}_ep$N.json 2> gpurun_out/bench_r02_${w}_ep$N.err
  echo "bench $w ep$N rc=$?"; tail -1 gpurun_out/bench_r02_${w}_ep$N.json | cut -c1-330; grep "ep phases" gpurun_out/bench_r02_${w}_ep$N.err | head -8; grep -i "error\|Traceback" -A5 gpurun_out/bench_r02_${w}_ep$N.err | head -20
done
