#!/bin/bash
# library with the re-tuned planner of the tile-resident kernel: whole GPU suite, smoke, small-lattice bench, bench.py
set -x
timeout 1200 python -m pytest tests -m gpu -x -q -p no:cacheprovider 2>&1 | tail -3
timeout 60 python __graft_entry__.py smoke 2>&1 | tail -1
timeout 300 python scripts/bench_small.py 2>&1 | grep "^{" | tee gpurun_out/r2zi_bench_small.jsonl | cut -c1-260
timeout 900 python bench.py > gpurun_out/r2zi_bench_n1.json 2> gpurun_out/r2zi_bench_n1.err; tail -c 200 gpurun_out/r2zi_bench_n1.err
