#!/bin/bash
# bench.py under torchrun exactly as the driver launches it (N ranks, --steps 20 --warmup 5), then the same with the plain
# mapping step (all-reduce + replicated FusedAdam) for an A/B of the fused reduce-scatter + Adam + all-gather kernel.
N=${1:-2}
mkdir -p gpurun_out
for step in fused plain; do
  extra=""; [ $step = plain ] && extra="--mapping-step plain --no-e2e"
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 $extra > gpurun_out/bench_n${N}_$step.json 2> gpurun_out/bench_n${N}_$step.err; echo "$step rc=$?"
  python - <<P
import json
try:
    d=json.loads(open('gpurun_out/bench_n${N}_$step.json').read().strip().splitlines()[-1])
    m=d.get('mapping') or {}
    print('N=$N $step value %.1f ms %.4f' % (d['value'], d['ms_per_step']), 'e2e', d['e2e'] and round(d['e2e']['value'],1), '| mapping frames/s %.2f ms %.4f' % (m.get('frames_per_s'), m.get('ms_per_step')))
    print('    ', m.get('collective'), '|', m.get('optimizer'))
    print('    ', m.get('collective_check'))
except Exception as e:
    print('unparsable', e); print(open('gpurun_out/bench_n${N}_$step.err').read()[-1500:])
P
done
