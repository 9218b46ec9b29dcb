#!/bin/bash
# N-GPU visit: fused mapping iteration (peer collective) + headline bench at N GPUs.
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519"
timeout 300 $TR bench_slam.py --mode mapping --impl ours --optimizer fused --gpus $N --collective peer --steps 96 > gpurun_out/slam_mapping_n${N}_peer.json 2> gpurun_out/slam_mapping_n${N}_peer.err; echo "slam mapping n$N rc=$?"
cut -c1-700 gpurun_out/slam_mapping_n${N}_peer.json
timeout 300 $TR bench.py --gpus $N --no-cpu-baseline --no-e2e --collective peer > gpurun_out/bench_n${N}_peer.json 2> gpurun_out/bench_n${N}_peer.err; echo "bench n$N rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_n${N}_peer.json')); print('bench value', d['value'], 'ms/step', d['ms_per_step'], d['config']['parallelism'])"
