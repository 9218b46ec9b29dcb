#!/bin/bash
# GPU suite (both record-feed modes), gather A/B on the headline bench, fused tracking / mapping iteration kernels.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
SPLAT_B200_GATHER=1 timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_slam_glue.py -m gpu -q -x > gpurun_out/pytest_gpu_gather.log 2>&1; echo "pytest gather rc=$?" >> gpurun_out/pytest_gpu_gather.log
tail -3 gpurun_out/pytest_gpu_gather.log
for g in 0 1 0 1; do
  SPLAT_B200_GATHER=$g timeout 600 python bench.py --no-cpu-baseline --no-e2e --no-mapping --warmup 30 > gpurun_out/ab_gather$g.json 2> gpurun_out/ab_gather$g.err || { echo "gather=$g failed"; tail -3 gpurun_out/ab_gather$g.err; continue; }
  python -c "
import json; d=json.loads(open('gpurun_out/ab_gather$g.json').read().strip().splitlines()[-1])
print('gather=$g', 'value %.1f ms %.4f' % (d['value'], d['ms_per_step']), d['kernel_ms_per_step'])"
done
for g in 0 1; do for mode in tracking mapping; do
  extra=""; [ $mode = tracking ] && extra="--pose-only"
  SPLAT_B200_GATHER=$g timeout 300 python bench_slam.py --mode $mode --impl ours --optimizer fused $extra --steps 96 --warmup 30 > gpurun_out/slam_${mode}_g$g.json 2> gpurun_out/slam_${mode}_g$g.err; echo "$mode rc=$?"
  python -c "
import json; d=json.loads(open('gpurun_out/slam_${mode}_g$g.json').read().strip().splitlines()[-1])
print('$mode gather=$g value %.1f ms %.4f' % (d['value'], d['ms_per_step'])); print(d.get('kernel_ms_per_step'))"
done; done
cat gpurun_out/tie_sort_time.json | python -c "import json,sys; d=json.load(sys.stdin); print('tile_sort ties %.4f smooth %.4f' % (d['kernel_ms']['ties']['tile_sort'], d['kernel_ms']['smooth']['tile_sort']))"
