#!/bin/bash
# round 2, final single-GPU records: full GPU suite, bench lines of every workload, reference arm, c3 sweep, launch list
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -q -m gpu --timeout 900 > gpurun_out/pytest_gpu_r02.txt 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu_r02.txt
for w in c2 c1 c3 c4 c5; do
  timeout 900 python bench.py --workload $w --steps 20 --warmup 5 > gpurun_out/bench_r02_$w.json 2> gpurun_out/bench_r02_$w.err; echo "$w rc=$?"
  tail -1 gpurun_out/bench_r02_$w.json | cut -c1-170
done
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_r02_reference_arm.json 2> gpurun_out/bench_r02_reference_arm.err; echo "ref arm rc=$?"
: > gpurun_out/bench_r02_c3_bs_sweep.jsonl
for bs in 1 4 8 16 32 64; do MXB_BENCH_SUSTAINED=0 MXB_BENCH_COPY_ONLY=0 timeout 300 python bench.py --workload c3 --bs $bs --steps 50 --warmup 10 --no-cpu 2>/dev/null | tail -1 >> gpurun_out/bench_r02_c3_bs_sweep.jsonl; done; wc -l gpurun_out/bench_r02_c3_bs_sweep.jsonl
MXB_BENCH_SUSTAINED=0 MXB_BENCH_COPY_ONLY=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"gg_pair|hadamard|quant_rows|combine|permute" -c 400 --csv --log-file gpurun_out/launches_r02_c2.csv \
   python bench.py --workload c2 --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_launches.log 2>&1; echo "ncu rc=$?"
