#!/bin/bash
# Round-2 evidence on the final tree, 1 GPU: GPU suite, both bench arms (default flags), fused SLAM iterations (eager and
# graph-replayed), ncu launch list + --set full capture of the headline bench (each after a clean plain run of the same command).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench ours rc=$?"
timeout 900 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
for cfg in "tracking --pose-only" "tracking --pose-only --graph 40" "tracking --pose-only --graph 200" "mapping"; do
  tag=$(echo $cfg | tr -d ' -'); st=96; case "$cfg" in *graph*) st=400;; esac
  timeout 300 python bench_slam.py --mode $cfg --impl ours --optimizer fused --steps $st --warmup 30 > gpurun_out/slam_$tag.json 2> gpurun_out/slam_$tag.err; echo "$tag rc=$?"
done
python - <<'P'
import json,glob
for f in ['bench_ours','bench_ref']+[g[11:-5] for g in sorted(glob.glob('gpurun_out/slam_*.json')) if 'g0' not in g and 'g1' not in g]:
    try:
        d=json.loads(open('gpurun_out/%s.json'%f).read().strip().splitlines()[-1])
        print(f, 'value %.1f ms %.4f'%(d['value'], d['ms_per_step']), 'e2e', (d.get('e2e') or {}).get('value'), 'e2e_res', (d.get('e2e_resident_params') or {}).get('value'), 'frac', (d.get('roofline') or {}).get('frac'))
        if d.get('mapping'): print('    mapping frames/s', d['mapping']['frames_per_s'], 'ms', d['mapping']['ms_per_step'])
        if d.get('kernel_ms_per_step'): print('   ', d['kernel_ms_per_step'])
        if d.get('cpu_baseline'): print('    cpu', d['cpu_baseline']['value'], d['cpu_baseline']['cores'])
    except Exception as e: print(f, 'ERR', e)
P
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-mapping"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/plain.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'render_|preprocess_|tile_sort|scatter|tile_scan' -s 175 -c 7 -o gpurun_out/r02_prof $CMD > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log; ls -la gpurun_out/r02_prof.ncu-rep gpurun_out/r02_launches.csv
