#!/bin/bash
cd "$(dirname "$0")/../.."
for dl in "$@"; do
  python bench.py --no-next --no-cpu --no-e2e --steps 20 --warmup 5 --device-lanes $dl 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('device lanes $dl', round(d['ms_per_step'],4), round(d['value']), d['keypoints_per_step'], d['stereo_matches_per_step'])"
done
