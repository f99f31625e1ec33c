#!/bin/bash
# A/B of run-time tuning knobs of the extractor (no rebuild): frames/s and stage times per setting
run() { env $1 python bench.py --steps 100 --warmup 5 --no-match --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', round(d['value']), round(d['e2e']['value']), {k:round(v,4) for k,v in d['stage_ms_per_step'].items()})"; }
run "X=0"
run "CCM_ORB_NO_PDL=1"
run "CCM_ORB_CHUNKS=2"
