#!/bin/bash
# A/B of library builds under tools/var/ (lib_<name>.so) against the in-tree build: frames/s and stage times per build
L=motioncheck_ccm_slam_b200/libccm_orb_cuda.so
run() { python bench.py --steps 100 --warmup 5 --no-match --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['value']), round(d['e2e']['value']), {k:round(v,4) for k,v in d['stage_ms_per_step'].items()})"; }
run "in-tree"
cp $L /tmp/lib_orig.so
for f in tools/var/lib_*.so; do cp $f $L; run "$(basename $f .so)"; done
cp /tmp/lib_orig.so $L
