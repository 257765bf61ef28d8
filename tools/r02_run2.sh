#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_moe_ops.py tests/test_gpu_ref_kernel.py tests/test_gpu_layer.py tests/test_gpu_layer_configs.py -q --timeout 600 > gpurun_out/pytest_gpu_r02b.txt 2>&1; echo "pytest rc=$?"
tail -25 gpurun_out/pytest_gpu_r02b.txt
timeout 300 python tools/had_probe.py > gpurun_out/had_probe_r02b.txt 2>&1; cat gpurun_out/had_probe_r02b.txt
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/bench_r02b_c2.json 2> gpurun_out/bench_r02b_c2.err; echo "bench rc=$?"
cut -c1-300 gpurun_out/bench_r02b_c2.json; tail -5 gpurun_out/bench_r02b_c2.err
