#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log
run() { name=$1; shift; timeout 300 python bench_slam.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/$name.json").read().strip().splitlines()[-1]); print("  ", round(d["value"],1), "it/s", round(d["ms_per_step"],4), "ms wall", round(d["wall_ms_per_step"],4), d.get("kernel_ms_per_step"))
except Exception as e: print("  failed", e); print(open("gpurun_out/$name.err").read()[-2500:])
PY
}
run slam_tracking_ours_lite_fusedopt --mode tracking --impl ours --pose-only --optimizer fused --steps 200
run slam_tracking_ours_lite_graph40 --mode tracking --impl ours --pose-only --optimizer fused --graph 40 --steps 200
run slam_tracking_ours_lite_graph200 --mode tracking --impl ours --pose-only --optimizer fused --graph 200 --steps 400
run slam_tracking_ours_graph40 --mode tracking --impl ours --optimizer fused --graph 40 --steps 200
