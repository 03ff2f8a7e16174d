python -m pytest tests -x -q -m gpu 2>&1 | grep -E "^E  |passed|failed|Error|^FAILED" | head -12
for c in C3 C4; do
  python bench.py --config $c --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --path two-pass 2>&1 | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print('$c two-pass', 'fps', round(d['value']), 'frac', round(r['frac'],4), 'ms', round(d['ms_per_step'],3))"
done
