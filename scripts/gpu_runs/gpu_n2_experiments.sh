N=${N:-2}
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513"
B="bench.py --gpus $N --steps 3 --warmup 3 --inner 300 --no-e2e"
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["value"], d["roofline"]["avg_launch_us"], d["clocks"]["sm_mhz"])'
timeout 300 $R scripts/slab_check.py --res 64 2>&1 | grep slab_check | tail -3
for T in p2p p2p-direct nccl p2p p2p-direct; do echo "== $T"; LB_SLAB_TRANSPORT=$T timeout 300 $R $B 2>&1 | grep '^{' | python -c "$P"; done
