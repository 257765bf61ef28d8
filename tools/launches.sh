#!/bin/bash
# ncu launch list (per-launch gpu__time_duration) of OUR kernels for one bench workload.
#   tools/launches.sh <workload> <tag> [extra bench args]
# Writes gpurun_out/launches_<tag>.csv and a per-kernel summary gpurun_out/launches_<tag>.txt.  Run under gpurun.
W=$1; TAG=$2; shift 2
mkdir -p gpurun_out
ncu --metrics gpu__time_duration.sum --clock-control none \
    -k 'regex:gg_|stream|quant_rows|combine|permute|hadamard|silu' -c 120 --csv \
    --log-file gpurun_out/launches_${TAG}.csv python bench.py --workload $W --steps 2 --warmup 3 --no-cpu "$@" > gpurun_out/ncu_${TAG}.log 2>&1
python tools/launch_summary.py gpurun_out/launches_${TAG}.csv > gpurun_out/launches_${TAG}.txt
cat gpurun_out/launches_${TAG}.txt
