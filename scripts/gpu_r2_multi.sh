#!/bin/bash
# bench.py under torchrun exactly as the driver launches it: N ranks, --steps 20 --warmup 5.
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "rc=$?"
tail -3 gpurun_out/bench_n$N.err
python - <<P
import json
d=json.loads(open('gpurun_out/bench_n$N.json').read().strip().splitlines()[-1])
print('N=$N value %.1f ms %.4f' % (d['value'], d['ms_per_step']), 'e2e', d['e2e'] and round(d['e2e']['value'],1), d['clocks'])
print('  ', d['config']['parallelism'])
print('   mapping', d.get('mapping'))
print('   idle', d.get('idle_ms_before_kernel_per_step'))
P
