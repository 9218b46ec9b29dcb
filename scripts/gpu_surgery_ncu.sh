#!/bin/bash
# ncu --set full of the map-surgery kernels (after a clean plain run).
mkdir -p gpurun_out
CMD="python bench_surgery.py --steps 2 --warmup 1"
timeout 300 $CMD > gpurun_out/surgery_plain.json 2> gpurun_out/surgery_plain.err || { echo plain failed; tail -5 gpurun_out/surgery_plain.err; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'compact_rows|prune_flags' -s 2 -c 2 -o gpurun_out/surgery_prof_prune $CMD > gpurun_out/surgery_ncu_prune.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'err_hist|spawn' -s 5 -c 5 -o gpurun_out/surgery_prof_grow $CMD > gpurun_out/surgery_ncu_grow.log 2>&1
ls -la gpurun_out/*.ncu-rep
