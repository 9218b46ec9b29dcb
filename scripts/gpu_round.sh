#!/bin/bash
# One GPU-box visit: parity tests, both bench arms, ncu launch list + full capture of the blend kernels.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench ours rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
timeout 600 python bench.py --impl reference --variant hx --no-e2e > gpurun_out/bench_ref_hx.json 2> gpurun_out/bench_ref_hx.err
timeout 600 python bench.py --variant hx --no-e2e --no-cpu-baseline > gpurun_out/bench_ours_hx.json 2> gpurun_out/bench_ours_hx.err
cat gpurun_out/bench_ours.json gpurun_out/bench_ref.json gpurun_out/bench_ref_hx.json gpurun_out/bench_ours_hx.json
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'render_|preprocess_|tile_sort|scatter' -s 35 -c 9 -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
