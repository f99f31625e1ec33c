#!/bin/bash
# sweep of k_pyramid_fused's tile grid (CCM_ORB_PYR_TILES) and rows-per-item builds; prints the pyramid stage time
L=motioncheck_ccm_slam_b200/libccm_orb_cuda.so
cp $L /tmp/lib_orig.so
run() { python bench.py --steps 60 --warmup 5 --no-match --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['value']), d['stage_ms_per_step']['pyramid_resize'])"; }
for v in orig r1 r4; do
  [ $v != orig ] && cp tools/var/lib_$v.so $L
  for t in 4x8 4x6 3x8 2x8 4x12 6x8 2x16; do CCM_ORB_PYR_TILES=$t run "$v $t"; done
done
cp /tmp/lib_orig.so $L
CCM_ORB_NO_FUSED_PYRAMID=1 run "per-level"
