#!/bin/bash
# ncu visit (ONE ncu pass per box visit, after the same command ran clean without it):
#   gpu_ncu.sh list  -> per-launch durations of one short bench run (gpurun_out/launches.csv)
#   gpu_ncu.sh full  -> --set full capture of one step's worth of our kernels (gpurun_out/prof.ncu-rep)
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
if [ "$1" = "list" ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
  tail -3 gpurun_out/ncu_launches.log
else
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'render_|preprocess_|tile_sort|scatter|tile_scan|emit_records' -s 32 -c 8 -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
  tail -3 gpurun_out/ncu_full.log
fi
