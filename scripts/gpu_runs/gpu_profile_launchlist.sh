# launch list covering the timed region of the default bench command (short inner count)
CMD="python bench.py --steps 2 --warmup 3 --inner 10 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain_ll.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_r1_full.csv $CMD > gpurun_out/ncu_list2.log 2>&1
tail -2 gpurun_out/ncu_list2.log | cut -c1-300
