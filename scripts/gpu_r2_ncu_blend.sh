#!/bin/bash
# ncu --set full (with source counters) of the two blend kernels on the headline workload, after a clean plain run.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'render_(fwd|bwd)' -s 6 -c 2 -o gpurun_out/r02c_blend $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out/*.ncu-rep
