#!/bin/bash
# N-GPU A/B on bench.py's render leg (driver flags): exchange kernels with their barriers fused in vs separate one-warp
# barrier kernels, and (N >= 4) in-switch vs peer loads / stores.
N=${1:-8}
mkdir -p gpurun_out
run() {  # tag, env...
  tag=$1; shift
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29563 bench.py --gpus $N --steps 20 --warmup 5 --no-e2e --no-mapping --no-cpu-baseline > gpurun_out/barrier_ab_n${N}_$tag.json 2> gpurun_out/barrier_ab_n${N}_$tag.err
  python - <<P
import json
d=json.loads(open('gpurun_out/barrier_ab_n${N}_$tag.json').read().strip().splitlines()[-1])
print('$tag N=$N value %.1f ms %.4f' % (d['value'], d['ms_per_step']))
P
}
for rep in 1 2; do
  run fused$rep SPLAT_PEER_FUSED_BARRIER=1
  run split$rep SPLAT_PEER_FUSED_BARRIER=0
done
