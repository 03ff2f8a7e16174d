# more end-of-round evidence: full ncu captures of the streaming kernel (W720) and of the split-sample slab instance (RGB1),
# the ladder survey and the bwdif timing.  Each ncu run follows a plain run of the same command that exited 0.
set -x
for spec in "W720 stream stream 64" "RGB1 auto slab_scale 64" "C4 auto slab_scale 32"; do
  set -- $spec
  CMD="python bench.py --config $1 --path $2 --frames $4 --segments 1 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
  $CMD > gpurun_out/plain_$1.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:$3 -s 3 -c 1 -f -o gpurun_out/prof_$1 $CMD > gpurun_out/ncu_$1.log 2>&1
done
python tools/gpu_ladder.py > gpurun_out/ladder.txt 2>&1
python tools/gpu_bwdif_time.py > gpurun_out/bwdif_time.txt 2>&1
python tools/gpu_bwdif_time.py 720 576 600 yuv420p >> gpurun_out/bwdif_time.txt 2>&1
python tools/gpu_bwdif_time.py 1920 1080 120 yuv420p10le >> gpurun_out/bwdif_time.txt 2>&1
tail -3 gpurun_out/bwdif_time.txt
