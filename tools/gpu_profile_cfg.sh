# one ncu --set full capture of the dominant kernel of config $CFG (kernel-name regex $KRE), after a plain run exited 0
set -x
CMD="python bench.py --config ${CFG:-W720P} --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain_${CFG}.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:${KRE:-stream_scale} -s 3 -c 1 -f -o gpurun_out/prof_${CFG} $CMD > gpurun_out/ncu_${CFG}.log 2>&1
tail -1 gpurun_out/plain_${CFG}.log | cut -c1-300
