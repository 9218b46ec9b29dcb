#!/bin/bash
# Map surgery: parity tests, bench (ours vs reference flow), per-kernel launch list.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_densify.py -q -x 2>&1 | tail -3
timeout 600 python bench_surgery.py > gpurun_out/surgery.json 2> gpurun_out/surgery.err; echo "bench rc=$?"; tail -3 gpurun_out/surgery.err
python - <<'PY'
import json
for l in open('gpurun_out/surgery.json'):
    d = json.loads(l); print(d['metric'][:20], 'ours', d['value'], 'ref', d['reference_flow_ms'], 'x', d['speedup'], 'kernel_ms', d['kernel_ms'], 'frac', d['roofline']['frac'])
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'prune_flags|flag_count|block_scan|compact_rows|err_hist|spawn' -c 200 --csv --log-file gpurun_out/surgery_launches.csv python bench_surgery.py --steps 4 --warmup 1 > gpurun_out/ncu_surgery.log 2>&1
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open('gpurun_out/surgery_launches.csv')) if len(r) > 10 and r[0].isdigit()]
acc = collections.defaultdict(list)
for r in rows: acc[r[4].split('(')[0].replace('splat::','')].append(float(r[-1]))
for k, v in acc.items(): print(k, len(v), 'median ns', sorted(v)[len(v)//2])
PY
