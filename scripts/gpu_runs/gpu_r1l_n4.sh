R4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29520"
echo "V=2"; LB_VEC=2 timeout 300 $R4 scripts/bench_cfg5.py --mode weak --nz 128 --steps 40 2>&1 | grep '^{' | cut -c1-200
echo "V=4"; timeout 300 $R4 scripts/bench_cfg5.py --mode weak --nz 128 --steps 40 2>&1 | grep '^{' | cut -c1-200
echo "V=2 single"; LB_VEC=2 timeout 300 python scripts/bench_cfg5.py --mode weak --nz 128 --steps 40 2>&1 | grep '^{' | cut -c1-200
echo "V=2 nz=64"; LB_VEC=2 timeout 300 $R4 scripts/bench_cfg5.py --mode weak --nz 64 --steps 40 2>&1 | grep '^{' | cut -c1-200
