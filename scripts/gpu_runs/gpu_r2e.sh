#!/bin/bash
# round 2, call e (1 GPU): new tests (BASELINE sizes, VTK pipeline), small-kernel timings, reports inside a run,
# the halo path with the slab as its own neighbour (kernel-side cost of the transports), ncu of both
set -x
timeout 900 python -m pytest tests -m gpu -x -q -p no:cacheprovider 2>&1 | tail -6
timeout 300 python -m pytest tests/test_gpu_baseline_sizes.py -m gpu -x -q -s -p no:cacheprovider 2>&1 | grep -E "config|passed|failed" | tee gpurun_out/r2e_baseline_sizes.log
for m in none ll remote; do timeout 120 python scripts/bench_halo_self.py 256 400 $m 2>&1 | tail -1; done | tee gpurun_out/r2e_halo_self.jsonl
timeout 600 python scripts/bench_output.py 256 2>&1 | grep "^{" | tee gpurun_out/r2e_bench_output.jsonl
timeout 300 ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 30 -c 1 -o gpurun_out/r2e_ncu_halo_ll python scripts/bench_halo_self.py 256 40 ll > /dev/null 2>&1
ncu -i gpurun_out/r2e_ncu_halo_ll.ncu-rep --page raw --csv > gpurun_out/r2e_ncu_halo_ll_raw.csv 2>/dev/null
timeout 300 ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 30 -c 1 -o gpurun_out/r2e_ncu_plain python scripts/bench_halo_self.py 256 40 none > /dev/null 2>&1
ncu -i gpurun_out/r2e_ncu_plain.ncu-rep --page raw --csv > gpurun_out/r2e_ncu_plain_raw.csv 2>/dev/null
ls -la gpurun_out
