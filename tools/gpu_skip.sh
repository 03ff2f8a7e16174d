for m in 0 1 2 4 3 5 6 7; do
  TVS_DEBUG_SKIP=$m python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('skip=$m', 'ms/step', round(d['ms_per_step'],3), 'us/frame', round(d['ms_per_step']*1000/300,2))"
done
