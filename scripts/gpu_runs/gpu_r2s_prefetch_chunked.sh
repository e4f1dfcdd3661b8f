#!/bin/bash
# chunked prefetch kernel (default 2 tiles per CTA) as built for production + tile-resident kernel with 1024 threads
set -x
timeout 900 python -m pytest tests/test_gpu_round2.py -m gpu -x -q -p no:cacheprovider 2>&1 | tail -5
timeout 900 python scripts/bench_prefetch.py 2>&1 | grep "^{" | tee gpurun_out/r2s_bench_prefetch.jsonl | cut -c1-420
for k in 3 4 8; do
  LB_PF_CHUNK=$k timeout 300 python scripts/bench_prefetch.py --cfg4 2>&1 | grep "^{" | sed "s/^{/{\"LB_PF_CHUNK\": $k, /" | tee -a gpurun_out/r2s_bench_prefetch_cfg4_chunk.jsonl | cut -c1-330
done
timeout 300 python scripts/bench_small.py 2>&1 | grep "^{" | tee gpurun_out/r2s_bench_small.jsonl
for h in 4 6 8; do
  LB_RES_H=$h timeout 120 python scripts/bench_small.py --cfg1 2>&1 | grep "^{" | tee -a gpurun_out/r2s_bench_small_forced_h.jsonl
done
