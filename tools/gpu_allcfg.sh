for c in ${CFGS:-C1 C2 C3 C4 C5}; do
  python bench.py --config $c --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('$c', 'fps', round(d['value']), 'frac', round(r['frac'],4), 'GB/s', round(r['achieved']), 'path', r['kernel'], 'launches/step', r['launches_per_step'], 'ms', round(d['ms_per_step'],3))"
done
