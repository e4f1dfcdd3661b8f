R="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513"
timeout 300 $R scripts/slab_check.py --res 64 2>&1 | grep slab_check | grep -E "packed|OK"
P='import sys,json; d=json.loads(sys.stdin.read()); print(d["value"], d["roofline"]["avg_launch_us"])'
for v in 0 1 0 1; do
  if [ $v = 1 ]; then export LB_NO_PDL=1; else unset LB_NO_PDL; fi
  echo "LB_NO_PDL=$v"; timeout 300 $R bench.py --gpus 2 --steps 3 --warmup 3 --inner 500 --no-e2e 2>/dev/null | grep '^{' | python -c "$P"
done
