#!/bin/bash
# (1) tile-resident kernel for small 2-D lattices: parity against one launch per step / the loop kernel, timing with
#     the planner's h and with forced h; (2) the prefetch kernel with the dynamic tile scheduler (experiment, LB_PF_DYN)
set -x
timeout 600 python -m pytest tests/test_gpu_round2.py -m gpu -x -q -p no:cacheprovider -k "resident or prefetch" 2>&1 | tail -5
timeout 300 python scripts/bench_small.py 2>&1 | grep "^{" | tee gpurun_out/r2q_bench_small.jsonl
for h in 2 3 4 5 6 8; do
  LB_RES_H=$h timeout 120 python scripts/bench_small.py --cfg1 2>&1 | grep "^{" | tee -a gpurun_out/r2q_bench_small_forced_h.jsonl
done
LB_PF_DYN=1 timeout 300 python -m pytest tests/test_gpu_round2.py -m gpu -x -q -p no:cacheprovider -k "prefetch" 2>&1 | tail -3
LB_PF_DYN=1 timeout 600 python scripts/bench_prefetch.py --quick 2>&1 | grep "^{" | tee gpurun_out/r2q_bench_prefetch_dynamic.jsonl | cut -c1-300
# one node per thread (96 registers, 5 CTAs per SM) instead of two packed nodes (166 registers, 3 CTAs): classic column
LB_VEC=1 timeout 300 python scripts/bench_prefetch.py --quick 2>&1 | grep "^{" | head -3 | tee gpurun_out/r2q_bench_kbc_one_node_per_thread.jsonl | cut -c1-300
