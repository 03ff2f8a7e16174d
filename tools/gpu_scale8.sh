# every BASELINE config on all 8 GPUs of one box, and the C5 farm job (strong scaling) at 1 / 2 / 4 / 8; device-resident lines only
: > gpurun_out/scale8.jsonl
run() {  # $1 config, $2 ranks
  if [ "$2" = 1 ]; then python bench.py --config $1 --gpus 1 --steps 5 --warmup 3 --no-e2e --no-cpu-baseline 2>>gpurun_out/scale8.err | tail -1 >> gpurun_out/scale8.jsonl
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port $((29540 + $2)) bench.py --config $1 --gpus $2 --steps 5 --warmup 3 --no-e2e --no-cpu-baseline 2>>gpurun_out/scale8.err | tail -1 >> gpurun_out/scale8.jsonl; fi
}
for n in 1 2 4 8; do run C5 $n; done
for c in C1 C2 C3 C4; do run $c 8; done
python - <<'PY'
import json
for ln in open('gpurun_out/scale8.jsonl'):
    try: d = json.loads(ln)
    except Exception: print('bad', ln[:100]); continue
    print(d['config']['workload'].split(':')[0], 'gpus', d['n_gpus'], d['scaling'], 'fps', round(d['value']), 'ms/step', round(d['ms_per_step'], 2), 'frac(per GPU)', round(d['roofline']['frac'], 4))
PY
