#!/bin/bash
# round 2, single GPU: the no-allocation permute test alone in a fresh process, the full GPU suite (which ends with it),
# then the bench lines of every workload and the ncu launch list of the c2 command
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu --timeout 300 > gpurun_out/pytest_no_alloc_alone_r02.txt 2>&1; echo "no_alloc alone rc=$?"; tail -3 gpurun_out/pytest_no_alloc_alone_r02.txt
timeout 2400 python -m pytest tests -q -m gpu --timeout 900 > gpurun_out/pytest_gpu_r02.txt 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu_r02.txt
for w in c2 c1 c3 c4 c5; do
  timeout 900 python bench.py --workload $w --steps 20 --warmup 5 > gpurun_out/bench_r02_$w.json 2> gpurun_out/bench_r02_$w.err; echo "$w rc=$?"
  tail -1 gpurun_out/bench_r02_$w.json | cut -c1-400
done
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_r02_reference_arm.json 2> gpurun_out/bench_r02_reference_arm.err; echo "ref arm rc=$?"; tail -1 gpurun_out/bench_r02_reference_arm.json | cut -c1-300
MXB_BENCH_SUSTAINED=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r02_c2.csv \
   python bench.py --workload c2 --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_launches.log 2>&1; echo "ncu rc=$?"
