#!/bin/bash
# round 2, call d (2 GPUs): LL transport with the incoming-halo loads behind the population loads; A/B against the
# handshake transports at the same build
set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 scripts/slab_check.py --res 48 --steps 21 > gpurun_out/r2d_slab_check_n2_full.log 2>&1
grep "slab_check" gpurun_out/r2d_slab_check_n2_full.log | tee gpurun_out/r2d_slab_check_n2.log
for t in p2p-ll p2p-packed p2p; do
  LB_SLAB_TRANSPORT=$t timeout 300 $TR --master-port 29513 bench.py --gpus 2 --no-extra --no-e2e 2>/dev/null | tail -1 | tee gpurun_out/r2d_bench_n2_$t.json
done
