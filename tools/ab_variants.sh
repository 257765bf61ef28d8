#!/bin/bash
# one gpurun call: parity tests of the streaming path, then c3 with the fused silu*mul epilogue on / off, then per-CTA timelines
L=$PWD/mxmoe_b200/lib
(timeout 300 python -m pytest tests/test_gpu_group_gemm.py tests/test_gpu_layer.py tests/test_gpu_moe_ops.py tests/test_gpu_checkpoint.py tests/test_gpu_moe_block.py -x -q -m gpu -k "stream or weight_only or decode or permute or layer or checkpoint or block" 2>&1 | tail -5)
echo "== epi on";  timeout 100 bash tools/bench_ab.sh c3
echo "== epi off"; MXB_STREAM_EPI=0 timeout 100 bash tools/bench_ab.sh c3
echo "== profile epi on";  MXB_LIB=$L/libmxmoe_b200_prof.so timeout 100 python tools/ws_profile.py 64 > gpurun_out/ws_prof_epi1.txt 2>&1; grep "==" gpurun_out/ws_prof_epi1.txt
echo "== profile epi off"; MXB_STREAM_EPI=0 MXB_LIB=$L/libmxmoe_b200_prof.so timeout 100 python tools/ws_profile.py 64 > gpurun_out/ws_prof_epi0.txt 2>&1; grep "==" gpurun_out/ws_prof_epi0.txt
