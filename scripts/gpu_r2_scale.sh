#!/bin/bash
# The driver's scaling run: bench.py at N = 1, 2, 4, 8 back to back (--steps 20 --warmup 5), ours only.
mkdir -p gpurun_out
for N in ${@:-1 2 4 8}; do
  if [ $N = 1 ]; then
    timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29520+N)) bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
  fi
  echo "N=$N rc=$?"
  python - <<P
import json
try:
    d=json.loads(open('gpurun_out/scale_n$N.json').read().strip().splitlines()[-1])
    m=d.get('mapping') or {}
    print('N=$N value %.1f ms %.4f' % (d['value'], d['ms_per_step']), 'e2e', d['e2e'] and round(d['e2e']['value'],1), '| mapping frames/s', m.get('frames_per_s'), 'ms', m.get('ms_per_step'), m.get('collective'))
    print('    ', d['config'].get('keyframe_schedule'), '|', m.get('keyframe_schedule'), '| idle pre_fwd', (d.get('idle_ms_before_kernel_per_step') or {}).get('preprocess_fwd'))
    print('    ', m.get('collective_check'))
except Exception as e:
    print('N=$N unparsable', e)
P
done
