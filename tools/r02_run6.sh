#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/pytest_gpu_r02i.txt 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu_r02i.txt
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r02i_c2.json 2> gpurun_out/bench_r02i_c2.err; echo "bench rc=$?"; cut -c1-250 gpurun_out/bench_r02i_c2.json; tail -3 gpurun_out/bench_r02i_c2.err
timeout 900 python bench.py --workload c5 --steps 10 --warmup 3 --no-cpu > gpurun_out/bench_r02i_c5.json 2> gpurun_out/bench_r02i_c5.err; echo "c5 rc=$?"; cut -c1-250 gpurun_out/bench_r02i_c5.json; tail -3 gpurun_out/bench_r02i_c5.err
bash tools/launches.sh c2 r02i_c2 > /dev/null 2>&1; cat gpurun_out/launches_r02i_c2.txt
