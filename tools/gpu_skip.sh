CFG=${CFG:-C1}
for m in 0 1 2 4 3 5 6 7; do
  TVS_DEBUG_SKIP=$m python bench.py --config $CFG --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --path fused 2>&1 | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); n=d['config']['frames_per_segment']; print('$CFG skip=$m', 'ms/step', round(d['ms_per_step'],3), 'us/frame', round(d['ms_per_step']*1000/n,2))"
done
