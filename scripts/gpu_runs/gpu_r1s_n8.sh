mkdir -p gpurun_out
R8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519"
timeout 200 $R8 scripts/slab_check.py --res 64 --steps 9 2>&1 | grep slab_check | grep -E "packed|OK"
: > gpurun_out/r1_cfg5_n8_packed_transport.jsonl
timeout 300 $R8 scripts/bench_cfg5.py --mode strong --steps 30 --warmup 5 2>&1 | grep '^{' | tee -a gpurun_out/r1_cfg5_n8_packed_transport.jsonl | cut -c1-200
timeout 300 $R8 bench.py --gpus 8 --steps 3 --warmup 3 2>/dev/null | grep '^{' | tee gpurun_out/r1_scale_packed_transport_n8.json | cut -c1-250
