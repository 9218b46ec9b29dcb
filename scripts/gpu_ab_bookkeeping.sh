#!/bin/bash
# A/B on one box: fused bookkeeping kernels (seen / best pose / conditional gradient scale) vs the torch statements
timeout 600 python -m pytest tests/test_gpu_slam_glue.py tests/test_gpu_slam_loop.py -q -x 2>&1 | tail -3
for v in 0 1 0 1; do
  echo "SPLAT_TORCH_BOOKKEEPING=$v"
  SPLAT_TORCH_BOOKKEEPING=$v timeout 300 python bench_slam.py --mode tracking --impl ours --optimizer fused --pose-only --graph 200 --steps 1000 --warmup 8 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('  tracking graph200 ms/iter', d['ms_per_step'])"
  SPLAT_TORCH_BOOKKEEPING=$v timeout 300 python bench_slam.py --mode mapping --impl ours --optimizer fused --steps 200 --warmup 8 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('  mapping eager ms/iter', d['ms_per_step'])"
done
