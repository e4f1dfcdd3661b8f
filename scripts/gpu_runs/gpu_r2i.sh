#!/bin/bash
# round 2, call i (1 GPU, DEVELOPMENT library with the D2Q9 units only): the chain kernel on config 1 and relatives
set -x
timeout 300 python scripts/bench_small.py 2>&1 | grep "^{" | tee gpurun_out/r2i_chain.jsonl
