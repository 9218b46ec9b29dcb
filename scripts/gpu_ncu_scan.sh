#!/bin/bash
# ncu --set full of the single-CTA tile_scan kernel (where do its 15 us go?)
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tile_scan -s 6 -c 2 -o gpurun_out/prof_scan $CMD > gpurun_out/ncu_scan.log 2>&1
ls -la gpurun_out | tail -5
