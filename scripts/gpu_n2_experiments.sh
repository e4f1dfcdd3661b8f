R="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513"
B="bench.py --gpus 2 --steps 3 --warmup 3 --inner 300 --no-e2e"
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["value"], d["roofline"]["avg_launch_us"])'
echo "== p2p (default)"; timeout 300 $R $B 2>&1 | grep '^{' | python -c "$P"
echo "== p2p, handshake kernels skipped (timing only, results invalid)"; LB_DEBUG_HALO_NOSYNC=1 timeout 300 $R $B 2>&1 | grep '^{' | python -c "$P"
echo "== nccl transport"; LB_SLAB_TRANSPORT=nccl timeout 300 $R $B 2>&1 | grep '^{' | python -c "$P"
