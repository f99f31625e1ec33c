#!/bin/bash
# A/B of the pyramid stage: per-level loads-first kernel (default), the older interleaved kernel, builds with other occupancy targets
L=motioncheck_ccm_slam_b200/libccm_orb_cuda.so
run() { python bench.py --steps 100 --warmup 5 --no-match --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['value']), round(d['e2e']['value']), d['stage_ms_per_step']['pyramid_resize'])"; }
run "per-level loads-first"
CCM_ORB_NO_RESIZE_MLP=1 run "per-level interleaved"
cp $L /tmp/lib_orig.so
for v in c5 c6; do [ -f tools/var/lib_$v.so ] && cp tools/var/lib_$v.so $L && run "loads-first $v"; done
cp /tmp/lib_orig.so $L
