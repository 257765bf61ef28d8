#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_group_gemm.py tests/test_gpu_layer.py tests/test_gpu_layer_configs.py -q -m gpu --timeout 600 -x > gpurun_out/pytest_r02m.txt 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_r02m.txt | cut -c1-250
for w in c2 c4 c5 c1; do MXB_BENCH_SUSTAINED=0 MXB_BENCH_COPY_ONLY=0 timeout 600 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu > gpurun_out/bench_r02m_$w.json 2> gpurun_out/bench_r02m_$w.err; echo "$w rc=$?"; python - <<PY
import json
d=json.loads(open("gpurun_out/bench_r02m_$w.json").read().strip().splitlines()[-1])
print("$w", round(d["ms_per_step"],4), "ms", round(d["value"],1), "gemm pair ms", round(d["roofline"]["ms_per_launch_pair"],4), "frac", round(d["roofline"]["frac"],3))
PY
done
