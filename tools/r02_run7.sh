#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_group_gemm.py tests/test_gpu_layer.py tests/test_gpu_ref_kernel.py tests/test_gpu_moe_block.py tests/test_gpu_checkpoint.py -q --timeout 900 > gpurun_out/pytest_gpu_r02j.txt 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu_r02j.txt
for w in c2 c4 c1; do timeout 600 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu > gpurun_out/bench_r02j_$w.json 2> gpurun_out/bench_r02j_$w.err; echo "$w rc=$?"; python - <<PY
import json
d=json.loads(open("gpurun_out/bench_r02j_$w.json").read().strip().splitlines()[-1])
print("$w", round(d["ms_per_step"],4), "ms", round(d["value"],1), "gemm pair ms", round(d["roofline"]["ms_per_launch_pair"],4), "frac", round(d["roofline"]["frac"],3))
PY
done
