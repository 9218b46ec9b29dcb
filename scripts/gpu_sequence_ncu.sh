#!/bin/bash
# per-kernel times of the C5 loop (short run) from an ncu launch list
mkdir -p gpurun_out
CMD="python bench_sequence.py --impl ours --frames 4 --sequence-frames 500 --tracking-iters 24 --mapping-iters 4"
timeout 300 $CMD > gpurun_out/seq_plain.json 2> gpurun_out/seq_plain.err || { echo plain failed; tail -5 gpurun_out/seq_plain.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/seq_launches.csv $CMD > gpurun_out/seq_ncu.log 2>&1
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open('gpurun_out/seq_launches.csv')) if len(r) > 10 and r[0].isdigit()]
acc = collections.defaultdict(list)
for r in rows: acc[r[4].split('(')[0].replace('splat::','')[:70]].append(float(r[-1]))
tot = sum(sum(v) for v in acc.values())
print('launches', len(rows), 'total ms', tot/1e6)
for k, v in sorted(acc.items(), key=lambda kv: -sum(kv[1]))[:40]: print('%-72s n=%5d med %8.0f ns sum %8.2f ms' % (k, len(v), sorted(v)[len(v)//2], sum(v)/1e6))
PY
