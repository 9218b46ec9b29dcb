#!/bin/bash
# A/B on one box: fused bookkeeping kernels vs the torch statements (graph-replayed tracking; short C5 run)
for v in 0 1 0 1; do
  echo "SPLAT_TORCH_BOOKKEEPING=$v"
  SPLAT_TORCH_BOOKKEEPING=$v timeout 300 python bench_slam.py --mode tracking --impl ours --optimizer fused --pose-only --graph 200 --steps 400 --warmup 8 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('  graph200 ms/iter', d['ms_per_step'])"
done
for v in 0 1; do
  SPLAT_TORCH_BOOKKEEPING=$v timeout 300 python bench_sequence.py --impl ours --frames 40 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('  C5 40 frames torch_bookkeeping=$v', d['value'], d['ms_per_frame_split'])"
done
