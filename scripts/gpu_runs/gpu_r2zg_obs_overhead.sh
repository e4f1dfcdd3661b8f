#!/bin/bash
# cost of the fused every-step reporters on the big lattices; obs_finish_kernel with programmatic dependent launch
set -x
timeout 1200 python -m pytest tests -m gpu -x -q -p no:cacheprovider 2>&1 | tail -3
timeout 600 python scripts/bench_obs_overhead.py 2>&1 | grep "^{" | tee gpurun_out/r2zg_obs_overhead_pdl.jsonl
