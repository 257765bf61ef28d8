#!/bin/bash
# wide-row Hadamard on the pipeline: parity tests, timing against the one-CTA-per-row kernel, c5 bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_moe_ops.py tests/test_gpu_layer_configs.py -q -m gpu --timeout 600 -k "hadamard or c5 or long_row" > gpurun_out/pytest_had_wide.txt 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_had_wide.txt
timeout 300 python tools/had_time.py 2>&1 | tee gpurun_out/had_time_r02.txt
MXB_HAD_PIPE_WIDE=0 timeout 300 python tools/had_time.py 8192 20480 2>&1 | tee -a gpurun_out/had_time_r02.txt
MXB_HAD_TRACE=1 timeout 120 python tools/had_one.py 8192 20480 2>&1 | tail -5 | cut -c1-1500 > gpurun_out/had_trace_wide.txt
timeout 900 python bench.py --workload c5 --steps 20 --warmup 5 --no-cpu > gpurun_out/bench_r02b_c5.json 2> gpurun_out/bench_r02b_c5.err; echo "c5 rc=$?"; tail -1 gpurun_out/bench_r02b_c5.json | cut -c1-200
