#!/bin/bash
# DEV library (five instantiation units).  (1) prefetch kernel with K consecutive tiles per CTA (hardware-scheduled)
# against persistent + dynamic counter; (2) tile-resident kernel with two nodes per trip, h sweep
set -x
for k in 2 4 8 16; do
  echo "LB_PF_CHUNK=$k"
  LB_PF_CHUNK=$k timeout 300 python scripts/bench_prefetch.py --quick 2>&1 | grep "^{" | head -4 | sed "s/^{/{\"LB_PF_CHUNK\": $k, /" | tee -a gpurun_out/r2r_bench_prefetch_chunk.jsonl | cut -c1-330
done
LB_PF_DYN=1 timeout 300 python scripts/bench_prefetch.py --quick 2>&1 | grep "^{" | head -4 | sed "s/^{/{\"LB_PF_DYN\": 1, /" | tee -a gpurun_out/r2r_bench_prefetch_chunk.jsonl | cut -c1-330
timeout 300 python scripts/bench_small.py 2>&1 | grep "^{" | tee gpurun_out/r2r_bench_small.jsonl
for h in 3 4 6 8 10; do
  LB_RES_H=$h timeout 120 python scripts/bench_small.py --cfg1 2>&1 | grep "^{" | tee -a gpurun_out/r2r_bench_small_forced_h.jsonl
done
