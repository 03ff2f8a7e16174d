# round-2 captures: C1 on the slab kernel at the bench's 300 frames (DRAM traffic per launch), C4 on the slab kernel
set -x
CMD="python bench.py --segments 1 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain_c1.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:slab_scale -s 3 -c 1 -f -o gpurun_out/prof_c1_300 $CMD > gpurun_out/ncu_c1_300.log 2>&1
CMD="python bench.py --config C4 --segments 1 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain_c4.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:slab_scale -s 3 -c 1 -f -o gpurun_out/prof_c4 $CMD > gpurun_out/ncu_c4.log 2>&1
tail -1 gpurun_out/plain_c4.log | cut -c1-200
