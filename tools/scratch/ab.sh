#!/bin/bash
# A/B of library variants on the GPU box: tools/scratch/ab.sh name1 name2 ... (snake-slam_b200/variants/<name>.so)
# AB_TEST=1 also runs the extractor parity tests with each variant.
cd "$(dirname "$0")/../.."
cp snake-slam_b200/libsnakeorb.so /tmp/keep.so
for v in "$@"; do
  cp snake-slam_b200/variants/$v.so snake-slam_b200/libsnakeorb.so
  if [ -n "$AB_TEST" ]; then echo "$v: $(timeout 300 python -m pytest tests/test_gpu_extract.py tests/test_gpu_fullsize.py -m gpu -x -q 2>&1 | tail -1)"; fi
  for rep in 1 2; do
    python bench.py --no-next --no-cpu --no-e2e --steps 20 --warmup 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d['roofline']['stage_ms_per_step']; print('$v', round(d['ms_per_step'],4), {k: round(x,4) for k,x in s.items()})"
  done
done
cp /tmp/keep.so snake-slam_b200/libsnakeorb.so
