#!/bin/bash
# kernel sequence of ONE graph-replayed tracking iteration (ncu launch list)
mkdir -p gpurun_out
CMD="python bench_slam.py --mode tracking --impl ours --optimizer fused --pose-only --graph 6 --steps 12 --warmup 2"
timeout 300 $CMD > /dev/null 2>&1 || { echo plain failed; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/iter_launches.csv $CMD > gpurun_out/iter_ncu.log 2>&1
python - <<'PY'
import csv
rows = [r for r in csv.reader(open('gpurun_out/iter_launches.csv')) if len(r) > 10 and r[0].isdigit()]
names = [r[4].split('(')[0].replace('splat::','')[:60] for r in rows]
dur = [float(r[-1]) for r in rows]
# last iteration = from the last rendervar_fwd to the end
idx = [i for i, n in enumerate(names) if n.startswith('rendervar_fwd')]
a, b = idx[-2], idx[-1]
print('launches per iteration:', b - a, ' sum us: %.1f' % (sum(dur[a:b]) / 1e3))
for i in range(a, b): print('%-62s %8.0f ns' % (names[i], dur[i]))
PY
