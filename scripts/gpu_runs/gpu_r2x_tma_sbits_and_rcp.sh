#!/bin/bash
# (1) TMA kernel with masked streaming + outlet (stock 2-D lt.Obstacle): parity, then bench.py (extra: config 3 stock list)
# (2) fp32 KBC on random fields: elements above 1e-5 with 1/feq = rcp.approx (as built) and with IEEE division
set -x
timeout 900 python -m pytest tests/test_gpu_round2.py tests/test_gpu_stock_flows.py tests/test_gpu_baseline_sizes.py -m gpu -x -q -p no:cacheprovider 2>&1 | tail -4
timeout 900 python bench.py > gpurun_out/r2x_bench_n1.json 2> gpurun_out/r2x_bench_n1.err; tail -c 300 gpurun_out/r2x_bench_n1.err
timeout 600 python scripts/kbc_rcp_count.py 2>&1 | grep "^{" | tee gpurun_out/r2x_kbc_rcp_count.jsonl | cut -c1-250
LETTUCE_B200_LIB=$PWD/scripts/dev/liblettuce_b200_exact_rcp.so timeout 600 python scripts/kbc_rcp_count.py 2>&1 | grep "^{" | tee -a gpurun_out/r2x_kbc_rcp_count.jsonl | cut -c1-250
