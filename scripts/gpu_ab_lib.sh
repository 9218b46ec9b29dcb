#!/bin/bash
# A/B of alternative builds of libsplat_b200.so on ONE box: usage gpu_ab_lib.sh tag=path [tag=path ...]
# ("default" = the in-tree library).  Prints value / kernel times of the headline bench per build.
mkdir -p gpurun_out
for spec in "$@"; do
  tag=${spec%%=*}; path=${spec#*=}
  if [ "$path" = "default" ]; then unset SPLAT_B200_LIB; else export SPLAT_B200_LIB=$PWD/$path; fi
  timeout 600 python bench.py --no-cpu-baseline --no-e2e --no-mapping --warmup 30 > gpurun_out/ab_$tag.json 2> gpurun_out/ab_$tag.err || { echo "$tag failed"; tail -3 gpurun_out/ab_$tag.err; continue; }
  python -c "
import json; d=json.loads(open('gpurun_out/ab_$tag.json').read().strip().splitlines()[-1])
print('$tag', 'value %.1f ms %.4f' % (d['value'], d['ms_per_step']), {k: round(v, 4) for k, v in d['kernel_ms_per_step'].items()})"
done
