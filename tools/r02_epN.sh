#!/bin/bash
# usage: tools/r02_epN.sh N "<pytest -k expr>" "<workloads>"   — EP parity tests at W = N, then bench lines at N GPUs
N=$1; K=$2; WL=$3
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 1200 python -m pytest tests/test_gpu_ep.py -q -m gpu --timeout 600 -k "$K" -rA > gpurun_out/pytest_ep_r02_n$N.txt 2>&1; echo "pytest rc=$?"; grep -E "PASSED|FAILED|SKIPPED|passed|failed" gpurun_out/pytest_ep_r02_n$N.txt | tail -12
for w in $WL; do
  extra=""; [ "$w" != "c2" ] && extra="MXB_BENCH_SUSTAINED=0"
  st=20; wu=5; [ "$w" = "c5_stack" ] && { st=5; wu=3; }
  env $extra MXB_EP_TIMING=1 timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --workload $w --steps $st --warmup $wu > gpurun_out/bench_r02_${w}_ep$N.json 2> gpurun_out/bench_r02_${w}_ep$N.err
  echo "bench $w ep$N rc=$?"; tail -1 gpurun_out/bench_r02_${w}_ep$N.json | cut -c1-160
  grep "ep phases" gpurun_out/bench_r02_${w}_ep$N.err > gpurun_out/ep_phases_r02_${w}_n$N.txt; head -3 gpurun_out/ep_phases_r02_${w}_n$N.txt | cut -c1-220
done
