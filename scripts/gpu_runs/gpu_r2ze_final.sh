#!/bin/bash
# library with the population base pointers passed by the host (prefetch kernel): whole GPU suite, smoke, bench.py,
# prefetch A/B, small-lattice bench
set -x
timeout 1200 python -m pytest tests -m gpu -x -q -p no:cacheprovider 2>&1 | tail -3
timeout 60 python __graft_entry__.py smoke 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/r2ze_bench_n1.json 2> gpurun_out/r2ze_bench_n1.err; tail -c 300 gpurun_out/r2ze_bench_n1.err
timeout 900 python scripts/bench_prefetch.py 2>&1 | grep "^{" | tee gpurun_out/r2ze_bench_prefetch.jsonl | cut -c1-100
timeout 600 python bench.py --impl reference > gpurun_out/r2ze_bench_reference.json 2>/dev/null; tail -c 400 gpurun_out/r2ze_bench_reference.json
timeout 300 python scripts/bench_small.py 2>&1 | grep "^{" | tee gpurun_out/r2ze_bench_small.jsonl | cut -c1-260
