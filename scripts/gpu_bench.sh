#!/bin/bash
# Quick GPU visit: parity tests + both bench arms (no ncu).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python bench.py ${BENCH_ARGS} > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench ours rc=$?"
timeout 900 python bench.py --impl reference ${BENCH_ARGS} > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
tail -n 3 gpurun_out/bench_ours.err; tail -n 3 gpurun_out/bench_ref.err
cat gpurun_out/bench_ours.json gpurun_out/bench_ref.json
