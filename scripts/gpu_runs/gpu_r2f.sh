#!/bin/bash
# round 2, call f (1 GPU): persistent TMA-prefetch kernel (parity + A/B), halo face phase (self-neighbour cost),
# vectorised enstrophy kernel, parallel file writes of the VTK pipeline
set -x
timeout 300 python -m pytest tests/test_gpu_round2.py -m gpu -x -q -p no:cacheprovider 2>&1 | tail -12
timeout 900 python -m pytest tests -m gpu -x -q -p no:cacheprovider 2>&1 | tail -6
for m in none ll remote; do timeout 120 python scripts/bench_halo_self.py 256 400 $m 2>&1 | tail -1; done | tee gpurun_out/r2f_halo_self.jsonl
timeout 900 python scripts/bench_tma.py 2>&1 | grep "^{" | tee gpurun_out/r2f_bench_tma.jsonl
timeout 600 python scripts/bench_output.py 256 2>&1 | grep "^{" | tee gpurun_out/r2f_bench_output.jsonl
