R2="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531"
timeout 300 $R2 scripts/slab_check.py 2>&1 | grep slab_check | tail -12
timeout 200 python -m pytest tests/test_multi_gpu.py -x -q 2>&1 | tail -3
timeout 300 $R2 bench.py --gpus 2 --steps 3 --warmup 3 2>/dev/null | tail -1 > gpurun_out/r1u_bench_n2.json; cut -c1-400 gpurun_out/r1u_bench_n2.json
