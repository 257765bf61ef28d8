#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 900 > gpurun_out/pytest_gpu_r02.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_r02.txt
cat > /tmp/had3.py <<'PY'
import os, sys, torch
sys.path.insert(0, os.getcwd())
from mxmoe_b200 import ops
h = ops.Handle(0)
for rows, n in ((8192, 20480), (32768, 1408), (8192, 5632), (65536, 2560)):
    x = (torch.rand(rows, n, device="cuda") * 4 - 2).half()
    q8 = torch.empty(rows, n, dtype=torch.int8, device="cuda"); sc = torch.empty(rows, 1, dtype=torch.float16, device="cuda")
    for _ in range(3):
        h.quant_act(x, None, rows, None, [(8, -1)], True, q8, None, sc)
    torch.cuda.synchronize()
    del x, q8, sc
PY
timeout 600 ncu --set full --clock-control none -k regex:hadamard_pipe -c 12 -o gpurun_out/ncu_had_pipe_all_r02 -f python /tmp/had3.py > gpurun_out/ncu_had3.log 2>&1; echo "ncu rc=$?"
