#!/bin/bash
# round 2, GPU call 1: every GPU test, the default bench line, the reference-kernel comparison, baseline of the rotation kernels
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/pytest_gpu_r02a.txt 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_r02a.txt
tail -15 gpurun_out/pytest_gpu_r02a.txt
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r02a_c2.json 2> gpurun_out/bench_r02a_c2.err; echo "bench rc=$?"
cut -c1-600 gpurun_out/bench_r02a_c2.json; tail -5 gpurun_out/bench_r02a_c2.err
timeout 600 python tools/ref_kernel_compare.py > gpurun_out/ref_kernel_compare_r02.jsonl 2> gpurun_out/ref_cmp.err; echo "refcmp rc=$?"; cat gpurun_out/ref_kernel_compare_r02.jsonl; tail -3 gpurun_out/ref_cmp.err
timeout 300 python tools/had_probe.py > gpurun_out/had_probe_r02a.txt 2>&1; cat gpurun_out/had_probe_r02a.txt
python tools/had_one.py 32768 1408 > gpurun_out/had_one.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:hadamard -s 2 -c 1 -o gpurun_out/had_umma11_r02a python tools/had_one.py 32768 1408 > gpurun_out/ncu_had.log 2>&1
echo "ncu rc=$?"
