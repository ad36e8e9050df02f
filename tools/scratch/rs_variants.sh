cd snake-slam_b200; cp libsnakeorb.so libsnakeorb_orig.so
for v in libsnakeorb_orig.so libsnakeorb_h*_u*.so; do
  cp $v libsnakeorb.so
  echo "== $v $(cd ..; timeout 300 python bench.py --no-cpu --no-next --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('value', round(d['value']), 'pyramid', round(d['roofline']['stage_ms_per_step']['pyramid'],4))")"
done
cp libsnakeorb_orig.so libsnakeorb.so
