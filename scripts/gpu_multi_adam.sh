#!/bin/bash
# N-GPU visit (gpurun --gpus N): peer all-reduce check + fused reduce/Adam check.
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513"
timeout 200 $TR tests/peer_allreduce_check.py > gpurun_out/peer_check_n$N.log 2>&1; echo "peer check rc=$?"
grep "PEER_ALLREDUCE\|fixed cost\|56 MB" gpurun_out/peer_check_n$N.log
timeout 200 $TR tests/peer_reduce_adam_check.py > gpurun_out/peer_adam_check_n$N.log 2>&1; echo "reduce+adam check rc=$?"
grep -v "^W\|^\*\|OMP_NUM\|NCCL version" gpurun_out/peer_adam_check_n$N.log | tail -8
