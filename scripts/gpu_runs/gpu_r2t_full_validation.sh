#!/bin/bash
# whole GPU suite, smoke, bench.py (production defaults: chunked prefetch kernel for D3Q27 KBC / TRT / Smagorinsky fp32,
# tile-resident kernel for small 2-D lattices), small-lattice bench
set -x
timeout 1200 python -m pytest tests -m gpu -x -q -p no:cacheprovider 2>&1 | tail -5
timeout 60 python __graft_entry__.py smoke 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/r2t_bench_n1.json 2> gpurun_out/r2t_bench_n1.err; tail -c 600 gpurun_out/r2t_bench_n1.err
timeout 300 python scripts/bench_small.py 2>&1 | grep "^{" | tee gpurun_out/r2t_bench_small.jsonl
