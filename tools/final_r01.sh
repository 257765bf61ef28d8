#!/bin/bash
# Round-end evidence run (one gpurun call): the full GPU suite, smoke(), every bench line, the launch lists and the
# ncu --set full capture of the streaming kernel.  Everything lands in gpurun_out/; the judged copies go to profiles/.
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -4) | tee gpurun_out/pytest_gpu.txt
(timeout 120 python __graft_entry__.py smoke 2>&1 | tail -2) | tee gpurun_out/smoke.txt
for w in c2 c1 c3 c4 c4_e4m3 c5; do
  timeout 300 python bench.py --workload $w > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; tail -c 400 gpurun_out/bench_$w.json | cut -c1-200
done
timeout 300 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
for w in c3 c2; do timeout 300 bash tools/launches.sh $w r01_$w > /dev/null 2>&1; done
timeout 300 ncu --set full --clock-control none --import-source on -k 'regex:w4a16_stream' -c 2 -s 6 -o gpurun_out/ws_full \
  python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_ws.log 2>&1
ncu -i gpurun_out/ws_full.ncu-rep --page raw --csv > gpurun_out/ws_full_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/ws_full_raw.csv > gpurun_out/ncu_w4a16_stream.txt 2>&1
head -12 gpurun_out/ncu_w4a16_stream.txt
