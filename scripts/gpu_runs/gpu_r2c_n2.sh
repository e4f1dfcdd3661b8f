#!/bin/bash
# round 2, call c (2 GPUs): slab correctness for every transport incl. the new handshake-free 'p2p-ll', stock obstacle in
# slabs (masks sharded), enstrophy across slabs; bench.py --gpus 2 with the default transport and, for A/B, round 1's 'p2p'
set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 scripts/slab_check.py --res 48 --steps 21 2>&1 | grep -v "^W\|^\*\*\*\|OMP_NUM" | tail -60 | tee gpurun_out/r2c_slab_check_n2.log
timeout 600 $TR --master-port 29512 bench.py --gpus 2 2>gpurun_out/r2c_bench_n2.err | tail -1 | tee gpurun_out/r2c_bench_n2.json
tail -5 gpurun_out/r2c_bench_n2.err
LB_SLAB_TRANSPORT=p2p timeout 300 $TR --master-port 29513 bench.py --gpus 2 --no-extra --no-e2e 2>&1 | tail -1 | tee gpurun_out/r2c_bench_n2_p2p.json
