#!/bin/bash
# Quick visit: full GPU test-suite, headline bench (no e2e / cpu legs), fused SLAM iteration in both modes.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-cpu-baseline --no-e2e > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench_quick.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step']); print(d.get('kernel_ms_per_step')); print(d.get('roofline',{}).get('frac'))"
for mode in tracking mapping; do
  timeout 300 python bench_slam.py --mode $mode --impl ours --steps 96 --warmup 8 > gpurun_out/slam_${mode}_ours.json 2> gpurun_out/slam_${mode}_ours.err; echo "$mode rc=$?"
  python -c "
import json; d=json.loads(open('gpurun_out/slam_${mode}_ours.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step']); print(d.get('kernel_ms_per_step'))"
done
