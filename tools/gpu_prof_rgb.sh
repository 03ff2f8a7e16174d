CMD="python bench.py --config RGB1 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain_RGB1.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum --clock-control none --csv --log-file gpurun_out/rgb_launches.csv -c 60 $CMD > gpurun_out/ncu_rgb.log 2>&1
tail -1 gpurun_out/plain_RGB1.log | cut -c1-200
