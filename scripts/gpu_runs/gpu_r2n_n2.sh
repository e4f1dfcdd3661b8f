#!/bin/bash
# 2 GPUs: slabs whose thickness is not a multiple of the vector width (one node per thread) against the undivided run
# (four nodes per thread) -- must be bit-exact after the packed-rounding fix
set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29511 scripts/slab_check.py --res 44 --steps 21 > gpurun_out/r2n_slab_check_n2_res44_full.log 2>&1
grep "slab_check" gpurun_out/r2n_slab_check_n2_res44_full.log | tee gpurun_out/r2n_slab_check_n2_res44.log | grep -v "bit-exact=True" 
