# one short device-resident bench line per config into gpurun_out/allcfg.jsonl (no CPU arm, no host copies)
: > gpurun_out/allcfg.jsonl
for c in ${CFGS:-C1 C2 C3 C4 C5 W1080 W720 W4K DVD480 HDR8 C4P8 C4S C4N RGB1 RGB2}; do
  python bench.py --config $c --steps ${STEPS:-10} --warmup 3 --no-cpu-baseline --no-e2e 2>gpurun_out/allcfg_err.log | tail -1 >> gpurun_out/allcfg.jsonl
done
python - <<'PY'
import json
for ln in open('gpurun_out/allcfg.jsonl'):
    try: d=json.loads(ln)
    except Exception: print('bad line', ln[:80]); continue
    r=d['roofline']; print(d['config']['workload'].split(':')[0], 'fps', round(d['value']), 'frac', round(r['frac'],4), 'fma_frac', r.get('fma_frac') and round(r['fma_frac'],3), r['kernel'], 'ms/seg', round(r['kernel_ms'],3), 'golden', d['golden_check']['frames'])
PY
