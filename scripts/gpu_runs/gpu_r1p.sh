ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 6 -c 1 -o gpurun_out/r1_sb_d3q19 -f python scripts/bench_configs.py --quick "--only=stock lt.Obstacle D3Q19" > gpurun_out/r1p_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 6 -c 1 -o gpurun_out/r1_sb_d2q9 -f python scripts/bench_configs.py --quick "--only=stock lt.Obstacle D2Q9" > gpurun_out/r1p_ncu2.log 2>&1
tail -3 gpurun_out/r1p_ncu2.log
