#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_hf_blocks.py -q -m gpu --timeout 600 > gpurun_out/pytest_hf_r02.txt 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_hf_r02.txt | cut -c1-250
MXB_BENCH_SUSTAINED=0 MXB_BENCH_COPY_ONLY=0 timeout 900 ncu --set full --import-source on --clock-control none -k regex:gg_pair -s 10 -c 2 -o gpurun_out/ncu_gg_pair_c2_r02 -f \
   python bench.py --workload c2 --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu_gg.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_gg.log | cut -c1-200
