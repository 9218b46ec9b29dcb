#!/bin/bash
# C5 like-for-like: both arms on the SAME 100 frames with the reference's keyframe_selection_overlap, then ours on 500.
mkdir -p gpurun_out
timeout 900 python bench_sequence.py --impl ours --frames 100 > gpurun_out/seq_ours_100.json 2> gpurun_out/seq_ours_100.err; echo "ours100 rc=$?"
timeout 1500 python bench_sequence.py --impl reference --frames 100 > gpurun_out/seq_ref_100.json 2> gpurun_out/seq_ref_100.err; echo "ref100 rc=$?"
timeout 900 python bench_sequence.py --impl ours --frames 500 > gpurun_out/seq_ours_500.json 2> gpurun_out/seq_ours_500.err; echo "ours500 rc=$?"
for f in seq_ours_100 seq_ref_100 seq_ours_500; do python -c "
import json; d=json.loads(open('gpurun_out/$f.json').read().strip().splitlines()[-1])
print('$f', 'frames/s %.3f' % d['value'], 'ms/frame %.1f' % d['ms_per_frame'], d['ms_per_frame_split'], 'ATE %.5f' % d['ate_rmse_m'], d['final_gaussians'], d['config']['keyframe_selection'][:40])" || tail -3 gpurun_out/$f.err; done
