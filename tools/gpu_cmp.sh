# frames/s of the configs in $CFGS on the kernel paths in $PATHS (short runs, no CPU arm, no host copies)
for p in ${PATHS:-fused slab}; do for c in ${CFGS:-C1 C2 W1080}; do
  python bench.py --config $c --path $p --segments ${SEGS:-4} --steps ${STEPS:-5} --warmup 3 --no-cpu-baseline --no-e2e 2>gpurun_out/cmp_err.log | tail -1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('$c $p', 'fps', round(d['value']), 'frac', round(d['roofline']['frac'],4), d['roofline']['kernel'], 'ms', round(d['ms_per_step'],3))
except Exception as e:
    print('$c $p failed', e)"
done; done
