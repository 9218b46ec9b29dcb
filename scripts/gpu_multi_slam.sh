#!/bin/bash
# N-GPU visit (gpurun --gpus N): the fused mapping iteration (get_loss + backward + all-reduce + FusedAdam), keyframe-sharded.
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
for c in peer nccl; do
  timeout 300 $TR bench_slam.py --mode mapping --impl ours --optimizer fused --gpus $N --collective $c --steps 96 > gpurun_out/slam_mapping_n${N}_$c.json 2> gpurun_out/slam_mapping_n${N}_$c.err; echo "slam mapping n$N $c rc=$?"
  tail -3 gpurun_out/slam_mapping_n${N}_$c.err | cut -c1-300
  cat gpurun_out/slam_mapping_n${N}_$c.json | cut -c1-900
done
