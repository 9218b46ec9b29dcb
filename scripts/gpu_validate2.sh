#!/bin/bash
# Round-end validation incl. the C5 sequence.
bash scripts/gpu_validate.sh
timeout 600 python bench_surgery.py > gpurun_out/surgery.json 2> gpurun_out/surgery.err; echo "surgery rc=$?"
timeout 900 python bench_sequence.py --impl ours > gpurun_out/seq_ours.json 2> gpurun_out/seq_ours.err; echo "seq ours rc=$?"; cut -c1-420 gpurun_out/seq_ours.json
