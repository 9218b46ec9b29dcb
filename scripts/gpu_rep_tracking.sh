for g in 1 0 1 0 1; do
  SPLAT_B200_GATHER=$g timeout 300 python bench_slam.py --mode tracking --impl ours --optimizer fused --pose-only --steps 96 --warmup 30 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('gather=$g value %.1f ms %.4f wall %.4f' % (d['value'], d['ms_per_step'], d['wall_ms_per_step']))"
done
