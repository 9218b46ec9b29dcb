#!/bin/bash
# Parity suite against the -DSPLAT_BOUNDS_CHECK build of the library (device-side traps on every index), both record-feed
# modes.  Stands in for compute-sanitizer (closed on this pool).  Build: see profiles/r02_bounds_check.txt.
mkdir -p gpurun_out
export SPLAT_B200_LIB=$PWD/scratch/libsplat_b200_boundscheck.so
for g in 1 0; do
  SPLAT_B200_GATHER=$g timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_slam_glue.py tests/test_gpu_densify.py tests/test_gpu_slam_loop.py -m gpu -q -x > gpurun_out/bounds_check_g$g.log 2>&1; echo "gather=$g rc=$?"
  tail -2 gpurun_out/bounds_check_g$g.log; grep -c "SPLAT_CHECK failed" gpurun_out/bounds_check_g$g.log
done
