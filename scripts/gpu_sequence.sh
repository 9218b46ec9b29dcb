#!/bin/bash
# C5: end-to-end TUM-shape sequence, both arms.
mkdir -p gpurun_out
timeout 900 python bench_sequence.py --impl ours > gpurun_out/seq_ours.json 2> gpurun_out/seq_ours.err; echo "ours rc=$?"; tail -3 gpurun_out/seq_ours.err; cut -c1-700 gpurun_out/seq_ours.json
timeout 600 python bench_sequence.py --impl reference > gpurun_out/seq_ref.json 2> gpurun_out/seq_ref.err; echo "ref rc=$?"; tail -3 gpurun_out/seq_ref.err; cut -c1-700 gpurun_out/seq_ref.json
