#!/bin/bash
# round 2, call b (1 GPU): full GPU suite with the LL transport / masked slabs / pruned instantiations, smoke, full bench line
set -x
timeout 400 python -m pytest tests -m gpu -x -q -p no:cacheprovider 2>&1 | tail -15
timeout 60 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 600 python bench.py 2>gpurun_out/r2b_bench_n1.err | tail -1 | tee gpurun_out/r2b_bench_n1.json
tail -5 gpurun_out/r2b_bench_n1.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 2>&1 | tail -1 | tee gpurun_out/r2b_bench_reference.json
