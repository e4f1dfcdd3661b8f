#!/bin/bash
# per-thread prefetch (cp.async) step kernel: parity against the classic kernel, then A/B timing on D3Q27
set -x
timeout 600 python -m pytest tests/test_gpu_round2.py -m gpu -x -q -p no:cacheprovider -k "prefetch" 2>&1 | tail -5
timeout 900 python scripts/bench_prefetch.py 2>&1 | grep "^{" | tee gpurun_out/r2p_bench_prefetch.jsonl | cut -c1-400
