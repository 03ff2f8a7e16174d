# one ncu --set full capture of the slab kernel on config $CFG, after a plain run exited 0
set -x
CMD="python bench.py --config ${CFG:-C1} --path ${KPATH:-slab} --frames ${FRAMES:-100} --segments 1 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain_slab.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:${KRE:-slab_scale} -s 3 -c 1 -f -o gpurun_out/prof_slab $CMD > gpurun_out/ncu_slab.log 2>&1
tail -1 gpurun_out/plain_slab.log | cut -c1-300
