#!/bin/bash
# A/B of the pyramid stage: per-level loads-first kernel (default), the older interleaved kernel, and the fused chain kernel
run() { python bench.py --steps 100 --warmup 5 --no-match --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['value']), round(d['e2e']['value']), d['stage_ms_per_step']['pyramid_resize'])"; }
run "per-level loads-first"
CCM_ORB_NO_RESIZE_MLP=1 run "per-level interleaved"
CCM_ORB_FUSED_PYRAMID=1 CCM_ORB_PYR_TILES=3x8 run "fused 3x8"
