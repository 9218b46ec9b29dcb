#!/bin/bash
# ncu --set full capture of ONE fused SplaTAM iteration's glue kernels + the blend backward (after a clean plain run).
mkdir -p gpurun_out
for mode in mapping tracking; do
  CMD="python bench_slam.py --mode $mode --impl ours --steps 2 --warmup 3"
  timeout 300 $CMD > gpurun_out/slam_plain_$mode.json 2> gpurun_out/slam_plain_$mode.err || { echo plain $mode failed; tail -5 gpurun_out/slam_plain_$mode.err; exit 1; }
  # matching launches: 24 rendervar_fwd at set-up, then 6 per step (builder fwd, loss x2-3, render_bwd, builder bwd[, pose_finalize])
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'rendervar|pose_finalize|loss|render_bwd' -s 42 -c 6 -o gpurun_out/slam_prof_$mode $CMD > gpurun_out/slam_ncu_full_$mode.log 2>&1
  tail -2 gpurun_out/slam_ncu_full_$mode.log
done
ls -la gpurun_out/*.ncu-rep
