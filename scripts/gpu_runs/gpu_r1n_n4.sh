mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_slab_emulated.py tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -5
R4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29520"
timeout 300 $R4 scripts/slab_check.py --res 64 2>&1 | grep slab_check | tail -9
for nz in 128 256 64; do for t in p2p-packed p2p-push; do
timeout 300 $R4 scripts/bench_cfg5.py --mode weak --nz $nz --steps 40 --transport $t 2>&1 | grep '^{' | cut -c1-170
done; done
timeout 300 $R4 scripts/bench_cfg5.py --mode weak --nxy 256 --nz 256 --steps 300 --transport p2p-packed 2>&1 | grep '^{' | cut -c1-170
timeout 300 $R4 scripts/bench_cfg5.py --mode weak --nxy 256 --nz 256 --steps 300 --transport p2p 2>&1 | grep '^{' | cut -c1-170
