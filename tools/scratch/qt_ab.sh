# quadtree distribution: big-level shared-memory capacity A/B (normal bench image and noisy variants)
for cfg in "0 2048" "2 8192"; do
  set -- $cfg
  echo "== big levels $1 cap $2"
  SNK_QT_BIG_LEVELS=$1 SNK_QT_BIG_CAP=$2 python bench.py --no-cpu --no-next --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('  bench: value', round(d['value']), 'stages', {k: round(v,3) for k,v in d['roofline']['stage_ms_per_step'].items()})"
  SNK_QT_BIG_LEVELS=$1 SNK_QT_BIG_CAP=$2 python tools/scratch/dense.py 2>/dev/null | sed 's/^/  /'
done
