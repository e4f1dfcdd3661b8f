python -m pytest tests/test_gpu_parity.py -m gpu -q -k "test_collisions_vs_oracle and kbc" 2>&1 | grep -E "^E  +assert|passed|failed" | head
CMD="python scripts/run_case.py D3Q27 kbc f32 256 12"
$CMD > gpurun_out/plain_kbc.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 6 -c 2 -o gpurun_out/prof_r1_d3q27_kbc_f32 $CMD > gpurun_out/ncu_kbc.log 2>&1
tail -2 gpurun_out/ncu_kbc.log
