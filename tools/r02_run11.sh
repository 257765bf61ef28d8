#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_moe_ops.py tests/test_gpu_layer_configs.py tests/test_gpu_hf_blocks.py tests/test_gpu_moe_block.py -q -m gpu --timeout 600 -k "hadamard or c5 or long_row or block" > gpurun_out/pytest_r02k.txt 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_r02k.txt | cut -c1-300
