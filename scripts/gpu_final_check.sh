#!/bin/bash
# What the driver runs at round end, on one GPU: the GPU suite, smoke(), and both bench arms with its flags.
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu 2>&1 | tail -1
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/drv_ours.json 2> gpurun_out/drv_ours.err; echo "ours rc=$?"
python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/drv_ref.json 2> gpurun_out/drv_ref.err; echo "reference rc=$?"
python - <<'P'
import json
for t in ("ours", "ref"):
    d = json.loads(open("gpurun_out/drv_%s.json" % t).read().strip().splitlines()[-1])
    print(t, "value %.1f ms %.4f" % (d["value"], d["ms_per_step"]), "e2e", (d.get("e2e") or {}).get("value"),
          "mapping frames/s", (d.get("mapping") or {}).get("frames_per_s"), "frac", (d.get("roofline") or {}).get("frac"),
          "launches", d.get("gpu_launches"), "cpu", (d.get("cpu_baseline") or {}).get("value"))
    if d.get("idle_ms_before_kernel_per_step"):
        print("    idle before kernels", d["idle_ms_before_kernel_per_step"])
P
