#!/bin/bash
# N-GPU A/B of the exchange kernels' barrier form (fused into the kernel vs separate one-warp kernels) on bench.py.
N=${1:-8}
mkdir -p gpurun_out
for fb in 1 0; do
  SPLAT_PEER_FUSED_BARRIER=$fb timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29560+fb)) bench.py --gpus $N --steps 20 --warmup 5 --no-e2e --no-cpu-baseline > gpurun_out/barrier_ab_n${N}_$fb.json 2> gpurun_out/barrier_ab_n${N}_$fb.err
  python - <<P
import json
d=json.loads(open('gpurun_out/barrier_ab_n${N}_$fb.json').read().strip().splitlines()[-1])
m=d.get('mapping') or {}
print('fused_barrier=$fb N=$N value %.1f ms %.4f' % (d['value'], d['ms_per_step']), '| mapping frames/s %.2f ms %.4f' % (m.get('frames_per_s', 0), m.get('ms_per_step', 0)), '| idle pre_fwd', d['idle_ms_before_kernel_per_step']['preprocess_fwd'])
P
done
