# end-of-step evidence: default bench line (all keys), reference arm, all-config table, ncu launch list and one full capture
set -x
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2>/dev/null
bash tools/gpu_allcfg.sh > gpurun_out/allcfg.txt 2>&1
CMD="python bench.py --segments 2 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain_short.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
CMD="python bench.py --frames 100 --segments 1 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain_short2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:slab_scale -s 3 -c 1 -f -o gpurun_out/prof_c1_full $CMD > gpurun_out/ncu_full.log 2>&1
cat gpurun_out/allcfg.txt
cut -c1-2500 gpurun_out/bench_default.json
