#!/bin/bash
# Round-end evidence: launch list + --set full capture of the headline bench, and --set full of one fused tracking
# (pose-only) and one fused mapping iteration incl. the Adam step.  Each ncu pass follows a clean plain run.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/plain.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'render_|preprocess_|tile_sort|scatter|tile_scan|emit_records' -s 32 -c 8 -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
CMD="python bench_slam.py --mode mapping --impl ours --optimizer fused --steps 2 --warmup 3"
timeout 300 $CMD > gpurun_out/slam_plain_mapping.json 2> gpurun_out/slam_plain_mapping.err || { echo plain mapping failed; exit 1; }
# matching launches: 24 rendervar_fwd at set-up, then 7 per step (builder fwd, loss x3, render_bwd, builder bwd, adam)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'rendervar|pose_finalize|loss|render_bwd|adam' -s 45 -c 7 -o gpurun_out/slam_prof_mapping $CMD > gpurun_out/slam_ncu_full_mapping.log 2>&1
tail -2 gpurun_out/slam_ncu_full_mapping.log
CMD="python bench_slam.py --mode tracking --impl ours --pose-only --optimizer fused --steps 2 --warmup 3"
timeout 300 $CMD > gpurun_out/slam_plain_tracking.json 2> gpurun_out/slam_plain_tracking.err || { echo plain tracking failed; exit 1; }
# per step: builder fwd, track_loss, loss_finalize, render_bwd, builder bwd, pose_finalize, adam = 7
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'rendervar|pose_finalize|loss|render_bwd|adam' -s 45 -c 7 -o gpurun_out/slam_prof_tracking $CMD > gpurun_out/slam_ncu_full_tracking.log 2>&1
tail -2 gpurun_out/slam_ncu_full_tracking.log
ls -la gpurun_out/*.ncu-rep gpurun_out/launches.csv
