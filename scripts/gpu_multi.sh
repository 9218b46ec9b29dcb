#!/bin/bash
# N-GPU visit (gpurun --gpus N): peer all-reduce check, then the bench with both collectives.
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513"
timeout 150 $TR tests/peer_allreduce_check.py > gpurun_out/peer_check_n$N.log 2>&1; echo "peer check rc=$?"
grep "PEER_ALLREDUCE\|fixed cost" gpurun_out/peer_check_n$N.log
for c in peer nccl; do
  timeout 300 $TR bench.py --gpus $N --no-cpu-baseline --collective $c > gpurun_out/bench_n${N}_$c.json 2> gpurun_out/bench_n${N}_$c.err; echo "bench $c rc=$?"
  python - <<PY
import json
d = json.load(open("gpurun_out/bench_n${N}_$c.json"))
print("$c", "value", d["value"], "ms/step", d["ms_per_step"], "e2e", d["e2e"]["value"] if d["e2e"] else None,
      "gap before preprocess_fwd", d["idle_ms_before_kernel_per_step"]["preprocess_fwd"], d["clocks"])
PY
done
