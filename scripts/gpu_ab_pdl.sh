#!/bin/bash
# A/B of programmatic dependent launch (SPLAT_B200_PDL=1/0) on one box: GPU suite with it on, headline bench, fused tracking
# iteration eager and graph-replayed, mapping iteration.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
for p in 1 0 1 0; do
  SPLAT_B200_PDL=$p timeout 600 python bench.py --no-cpu-baseline --no-e2e --no-mapping --warmup 30 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('pdl=$p bench value %.1f ms %.4f' % (d['value'], d['ms_per_step']), d['idle_ms_before_kernel_per_step'])"
done
for p in 1 0; do for cfg in "tracking --pose-only --graph 200" "tracking --pose-only --graph 40" "tracking --pose-only" "mapping"; do
  st=96; case "$cfg" in *graph*) st=400;; esac
  SPLAT_B200_PDL=$p timeout 300 python bench_slam.py --mode $cfg --impl ours --optimizer fused --steps $st --warmup 30 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('pdl=$p $cfg value %.1f ms %.4f' % (d['value'], d['ms_per_step']))"
done; done
