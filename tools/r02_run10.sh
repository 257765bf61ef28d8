#!/bin/bash
# Kronecker-widened Hadamard factor: parity tests, timing with and without, c5 bench, ncu source-level capture of the c2 kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_moe_ops.py tests/test_gpu_layer_configs.py tests/test_gpu_layer.py -q -m gpu --timeout 600 -k "hadamard or c5 or long_row or c2" > gpurun_out/pytest_had_kron.txt 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_had_kron.txt
timeout 300 python tools/had_time.py 2>&1 | tee gpurun_out/had_time_r02b.txt
for s in "65536 2560" "16384 14336" "8192 7168" "8192 10240"; do MXB_HAD_KRON=1 timeout 300 python tools/had_time.py $s 2>&1 | sed 's/$/ KRON=1/' | tee -a gpurun_out/had_time_r02b.txt; done
MXB_HAD_KRON=2 timeout 300 python tools/had_time.py 16384 14336 2>&1 | sed 's/$/ KRON=2/' | tee -a gpurun_out/had_time_r02b.txt
timeout 900 python bench.py --workload c5 --steps 20 --warmup 5 --no-cpu > gpurun_out/bench_r02c_c5.json 2> gpurun_out/bench_r02c_c5.err; echo "c5 rc=$?"; tail -1 gpurun_out/bench_r02c_c5.json | cut -c1-200
timeout 600 ncu --set full --import-source on --clock-control none -k regex:hadamard_pipe -s 2 -c 1 -o gpurun_out/ncu_had_pipe_c2_r02 -f python tools/had_one.py 32768 1408 > gpurun_out/ncu_had.log 2>&1; echo "ncu rc=$?"
