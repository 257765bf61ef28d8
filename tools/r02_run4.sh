#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_moe_ops.py tests/test_gpu_ref_kernel.py tests/test_gpu_layer.py -q --timeout 600 > gpurun_out/pytest_gpu_r02e.txt 2>&1; echo "pytest rc=$?"
tail -12 gpurun_out/pytest_gpu_r02e.txt
timeout 300 python tools/had_probe.py > gpurun_out/had_probe_r02e.txt 2>&1; cat gpurun_out/had_probe_r02e.txt
python tools/had_one.py 32768 1408 > gpurun_out/had_one.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:hadamard -s 2 -c 1 -o gpurun_out/had_pipe11_r02e python tools/had_one.py 32768 1408 > gpurun_out/ncu_had.log 2>&1
echo "ncu rc=$?"
