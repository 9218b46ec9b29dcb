#!/bin/bash
# Round-2 numbers on the non-headline workloads (ours; the reference arm's are unchanged since round 1).
mkdir -p gpurun_out
for wl in replica_1200x680_300k scannetpp_1752x1168_2M_hessian c1_64x64_10k; do
  timeout 600 python bench.py --workload $wl --no-e2e --no-cpu-baseline --no-mapping --steps 96 --warmup 30 > gpurun_out/wl_$wl.json 2> gpurun_out/wl_$wl.err; echo "$wl rc=$?"
done
timeout 600 python bench.py --rgbd fused --no-e2e --no-cpu-baseline --no-mapping --steps 96 --warmup 30 > gpurun_out/wl_rgbd_fused.json 2> gpurun_out/wl_rgbd_fused.err; echo "rgbd rc=$?"
timeout 600 python bench.py --variant dg --no-e2e --no-cpu-baseline --no-mapping --steps 96 --warmup 30 > gpurun_out/wl_dg.json 2> gpurun_out/wl_dg.err; echo "dg rc=$?"
timeout 600 python bench_surgery.py > gpurun_out/wl_surgery.json 2> gpurun_out/wl_surgery.err; echo "surgery rc=$?"
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/wl_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f[14:-5], 'value', d.get('value'), 'ms', d.get('ms_per_step'), {k:v for k,v in (d.get('kernel_ms_per_step') or {}).items() if 'render' in k})
    except Exception as e: print(f, 'ERR', e)
P
