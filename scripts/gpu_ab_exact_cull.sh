#!/bin/bash
# A/B on one box: exact per-sub-tile footprint test in BOTH blend kernels vs the bounding-box test in both.
# Build the variant first (here, not on the box):
#   mkdir -p scratch/csrc_exact && cp splatam_b200/csrc/*.cu splatam_b200/csrc/*.cuh splatam_b200/csrc/Makefile scratch/csrc_exact/
#   make -C scratch/csrc_exact -j8 EXTRA=-DSPLAT_EXACT_CULL=1 OUT=../libsplat_b200_exact.so
# (the shipped library uses the exact test in the backward only; -DSPLAT_EXACT_CULL=0 gives the old behaviour)
run() {
  timeout 300 python bench.py --steps 96 --warmup 8 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); k=d['kernel_ms_per_step']; print('  headline ms', round(d['ms_per_step'],4), 'fwd', k['render_fwd'], 'bwd', k['render_bwd'])"
  timeout 300 python bench_slam.py --mode tracking --impl ours --optimizer fused --pose-only --graph 200 --steps 600 --warmup 8 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('  tracking graph200 ms/iter', round(d['ms_per_step'],4))"
  timeout 300 python bench_sequence.py --impl ours --frames 30 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('  C5 30 frames', round(d['value'],3), d['ms_per_frame_split'])"
}
cp splatam_b200/libsplat_b200.so /tmp/lib_box.so
for v in box exact box exact; do
  if [ $v = exact ]; then cp scratch/libsplat_b200_exact.so splatam_b200/libsplat_b200.so; else cp /tmp/lib_box.so splatam_b200/libsplat_b200.so; fi
  echo "variant $v"; run
done
cp scratch/libsplat_b200_exact.so splatam_b200/libsplat_b200.so
timeout 600 python -m pytest tests/test_gpu_parity.py -q -x 2>&1 | tail -2
cp /tmp/lib_box.so splatam_b200/libsplat_b200.so
