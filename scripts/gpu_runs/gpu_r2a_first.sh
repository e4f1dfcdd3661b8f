#!/bin/bash
# First GPU call of the next round (prepared at the end of round 1, when the GPU budget had run out):
#   1. the whole GPU suite (567 tests; ~25 s)            2. smoke()
#   3. cost of one VTK report with lb_output_fields beside the torch route it replaced (expected ~0.24 ms vs 0.82 ms)
#   4. the driver-contract bench line                    5. ncu: launch list + one full capture of output_fields_kernel
# usage: gpurun --timeout 600 -- 'bash scripts/gpu_runs/gpu_r2a_first.sh > gpurun_out/r2a_first.log 2>&1'
set -x
timeout 200 python -m pytest tests -m gpu -x -q -p no:cacheprovider 2>&1 | tail -8
timeout 60 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 90 python scripts/bench_output.py 256 3 2>&1 | tail -2 | tee gpurun_out/r2_bench_output.json
timeout 200 python bench.py 2>&1 | tail -1 | tee gpurun_out/r2_bench_n1.json
timeout 120 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv \
    --log-file gpurun_out/r2_launches_bench_output.csv python scripts/bench_output.py 256 1 > /dev/null 2>&1
timeout 180 ncu --set full --clock-control none --import-source on -k regex:output_fields -c 1 \
    -o gpurun_out/r2_ncu_output_fields python scripts/bench_output.py 256 1 > /dev/null 2>&1
ncu -i gpurun_out/r2_ncu_output_fields.ncu-rep --page raw --csv > gpurun_out/r2_ncu_output_fields_raw.csv 2>/dev/null
