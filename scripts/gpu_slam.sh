#!/bin/bash
# One GPU-box visit for the SplaTAM-glue kernels: parity tests, then the whole-iteration bench in
# every arm.  Results under gpurun_out/.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_slam_glue.py -m gpu -q > gpurun_out/pytest_slam.log 2>&1; echo "pytest slam rc=$?" >> gpurun_out/pytest_slam.log
tail -40 gpurun_out/pytest_slam.log
for mode in tracking mapping; do
  for impl in ours dropin reference; do
    timeout 300 python bench_slam.py --mode $mode --impl $impl --steps 48 --warmup 6 > gpurun_out/slam_${mode}_${impl}.json 2> gpurun_out/slam_${mode}_${impl}.err; echo "$mode $impl rc=$?"
    cat gpurun_out/slam_${mode}_${impl}.json
  done
done
timeout 300 python bench_slam.py --mode tracking --impl reference --variant hx --steps 16 --warmup 3 > gpurun_out/slam_tracking_reference_hx.json 2> gpurun_out/slam_tracking_reference_hx.err
cat gpurun_out/slam_tracking_reference_hx.json
