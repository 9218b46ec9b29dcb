#!/bin/bash
# Round-end validation: smoke(), full GPU suite, both bench arms with every leg, fused SLAM iteration, per-kernel times.
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench ours rc=$?"
timeout 900 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
cat gpurun_out/bench_ours.json gpurun_out/bench_ref.json
for mode in tracking mapping; do
  timeout 300 python bench_slam.py --mode $mode --impl ours --steps 96 --warmup 8 > gpurun_out/slam_${mode}_ours.json 2> gpurun_out/slam_${mode}_ours.err; echo "$mode rc=$?"
  python -c "
import json; d=json.loads(open('gpurun_out/slam_${mode}_ours.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step']); print(d.get('kernel_ms_per_step'))"
done
timeout 300 python bench_slam.py --mode tracking --impl ours --optimizer fused --pose-only --graph 200 --steps 400 --warmup 8 > gpurun_out/slam_tracking_graph200.json 2> gpurun_out/slam_tracking_graph200.err; echo "graph rc=$?"; cat gpurun_out/slam_tracking_graph200.json
