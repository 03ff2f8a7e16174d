python -m pytest tests -x -q -m gpu 2>&1 | grep -E "^E  |passed|failed" | head -12
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | tail -1 > gpurun_out/bench_last.json
python -c "
import json; d=json.load(open('gpurun_out/bench_last.json')); print('fps', round(d['value']), 'frac', round(d['roofline']['frac'],4), 'ms', round(d['ms_per_step'],3), d['clocks'])"
