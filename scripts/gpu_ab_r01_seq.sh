#!/bin/bash
# A/B on one box: round-1 tree (scratch/r01, git worktree of 9f6ac3d) vs current tree on C5, with a per-kernel profile of
# an eager pose-only tracking iteration on the final state.
mkdir -p gpurun_out
N=${1:-300}
(cd scratch/r01 && timeout 900 python bench_sequence.py --impl ours --frames $N --profile-end 2>/dev/null) > gpurun_out/seq_r01_p.json; echo "r01 rc=$?"
timeout 900 python bench_sequence.py --impl ours --frames $N --profile-end > gpurun_out/seq_cur_p.json 2>/dev/null; echo "cur rc=$?"
for f in seq_r01_p seq_cur_p; do python - <<P
import json
d=json.loads(open('gpurun_out/$f.json').read().strip().splitlines()[-1])
print('$f', 'frames/s %.3f' % d['value'], d['ms_per_frame_split'], 'ATE %.5f' % d['ate_rmse_m'])
print('   ', d['tracking_ms_per_frame_by_50'])
print('   ', d.get('final_state_tracking_kernel_ms'))
print('   ', d.get('final_state'))
P
done
