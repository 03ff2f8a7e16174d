set -x
python -m pytest tests -x -q -m gpu 2>&1 | tail -2
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain_short.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain_short2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:fused_scale -s 3 -c 1 -o gpurun_out/prof_c1_full $CMD > gpurun_out/ncu_full.log 2>&1
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2>/dev/null
CFGS="C1 C2 C3 C4 C5 W1080 W720 W4K DVD480 HDR8 C4P8 RGB1" bash tools/gpu_allcfg.sh > gpurun_out/allcfg.txt 2>&1
cat gpurun_out/allcfg.txt
