R4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29520"
for nz in 128 256 64; do for t in p2p p2p-push; do
timeout 300 $R4 scripts/bench_cfg5.py --mode weak --nz $nz --steps 40 --transport $t 2>&1 | grep '^{' | cut -c1-200
done; done
timeout 300 $R4 scripts/bench_cfg5.py --mode weak --nxy 512 --nz 512 --steps 40 --transport p2p 2>&1 | grep '^{' | cut -c1-200
