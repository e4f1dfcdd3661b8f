R4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29520"
echo "nosync nz=128"; LB_DEBUG_HALO_NOSYNC=1 timeout 300 $R4 scripts/bench_cfg5.py --mode weak --nz 128 --steps 40 2>&1 | grep '^{' | cut -c1-160
for nz in 160 192 224 320; do echo "nz=$nz"; timeout 300 $R4 scripts/bench_cfg5.py --mode weak --nz $nz --steps 40 2>&1 | grep '^{' | cut -c1-160; done
echo "2048x512 nz=128"; timeout 300 $R4 scripts/bench_cfg5.py --mode weak --nxy 724 --nz 256 --steps 40 2>&1 | grep '^{' | cut -c1-160
