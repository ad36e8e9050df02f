#!/bin/bash
cd "$(dirname "$0")/../.."
for cfg in "$@"; do
  set -- $cfg
  python bench.py --no-next --no-cpu --steps 20 --warmup 5 --chunks $1 --lanes $2 --depth ${3:-2} 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('chunks $1 lanes $2 depth ${3:-2}', 'device', round(d['ms_per_step'],3), 'e2e ms', round(d['e2e']['ms_per_step'],3), round(d['e2e']['value']))"
done
