mkdir -p gpurun_out
nvidia-smi -L | wc -l
R8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519"
R4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29520"
timeout 300 $R8 scripts/slab_check.py --res 64 2>&1 | grep slab_check | tail -7
: > gpurun_out/cfg5_n8.jsonl
timeout 600 $R8 scripts/bench_cfg5.py --mode strong --steps 50 2>&1 | grep '^{' | tee -a gpurun_out/cfg5_n8.jsonl
timeout 600 $R4 scripts/bench_cfg5.py --mode strong --steps 50 2>&1 | grep '^{' | tee -a gpurun_out/cfg5_n8.jsonl
timeout 600 $R8 scripts/bench_cfg5.py --mode weak --steps 50 2>&1 | grep '^{' | tee -a gpurun_out/cfg5_n8.jsonl
timeout 600 $R4 scripts/bench_cfg5.py --mode weak --steps 50 2>&1 | grep '^{' | tee -a gpurun_out/cfg5_n8.jsonl
for n in 8 4 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --steps 3 --warmup 3 2>/dev/null | grep '^{' | tee gpurun_out/bench_n$n.json | cut -c1-200
done
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline | tee gpurun_out/bench_n1.json | cut -c1-200
