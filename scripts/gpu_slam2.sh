#!/bin/bash
# Full GPU test-suite + ncu evidence for the SplaTAM-glue kernels (one visit).
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
CMD="python bench_slam.py --mode mapping --impl ours --steps 2 --warmup 3"
timeout 300 $CMD > gpurun_out/slam_plain.json 2> gpurun_out/slam_plain.err || { echo plain failed; tail -5 gpurun_out/slam_plain.err; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/slam_launches.csv $CMD > gpurun_out/slam_ncu_launches.log 2>&1
tail -2 gpurun_out/slam_ncu_launches.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'rendervar|pose_finalize|loss' -c 12 -o gpurun_out/slam_prof $CMD > gpurun_out/slam_ncu_full.log 2>&1
tail -2 gpurun_out/slam_ncu_full.log
CMD="python bench_slam.py --mode tracking --impl ours --steps 2 --warmup 3"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'rendervar|pose_finalize|loss' -c 10 -o gpurun_out/slam_prof_track $CMD > gpurun_out/slam_ncu_full_track.log 2>&1
ls -la gpurun_out | tail -12
