#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
run() { name=$1; shift; timeout 300 python bench_slam.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/$name.json").read().strip().splitlines()[-1]); print("  ", round(d["value"],1), "it/s", round(d["ms_per_step"],4), "ms", d.get("kernel_ms_per_step"))
except Exception as e: print("  failed", e); print(open("gpurun_out/$name.err").read()[-1500:])
PY
}
run slam_tracking_ours --mode tracking --impl ours
run slam_tracking_ours_lite --mode tracking --impl ours --pose-only
run slam_tracking_ours_opt --mode tracking --impl ours --optimizer
run slam_tracking_ours_lite_opt --mode tracking --impl ours --pose-only --optimizer
run slam_tracking_reference_opt --mode tracking --impl reference --optimizer --steps 48
run slam_tracking_dropin_opt --mode tracking --impl dropin --optimizer --steps 48
run slam_mapping_ours_opt --mode mapping --impl ours --optimizer
run slam_mapping_reference_opt --mode mapping --impl reference --optimizer --steps 48
timeout 600 python bench.py --no-cpu-baseline --no-e2e > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench_quick.json').read().strip().splitlines()[-1])
print('value', d['value'], 'ms', d['ms_per_step']); print(d.get('kernel_ms_per_step'))"
