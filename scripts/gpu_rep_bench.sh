#!/bin/bash
# Run-to-run spread of the driver's N=1 command (20 timed steps): both legs, repeats with a per-step trace of the mapping leg.
for i in 1 2 3 4 5 6; do
  SPLAT_BENCH_TRACE=1 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline --no-e2e 2> gpurun_out/rep_$i.err | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); m=d['mapping']
print('run $i value %.1f ms %.4f | mapping ms %.4f frames/s %.2f' % (d['value'], d['ms_per_step'], m['ms_per_step'], m['frames_per_s']))"
  grep "per-step" gpurun_out/rep_$i.err
done
