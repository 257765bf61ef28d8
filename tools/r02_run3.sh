#!/bin/bash
mkdir -p gpurun_out
python tools/had_one.py 32768 1408 > gpurun_out/had_one.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:hadamard -s 2 -c 1 -o gpurun_out/had_pipe11_r02c python tools/had_one.py 32768 1408 > gpurun_out/ncu_had.log 2>&1
echo "ncu rc=$?"
