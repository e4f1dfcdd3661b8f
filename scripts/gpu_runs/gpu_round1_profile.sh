python -m pytest tests -m gpu -q 2>&1 | tail -40
CMD="python bench.py --steps 1 --warmup 3 --inner 20 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1.csv $CMD > gpurun_out/ncu_list.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 30 -c 3 -o gpurun_out/prof_r1_d3q19_bgk_f32 $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_list.log gpurun_out/ncu_full.log
