mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519"
timeout 600 $R scripts/bench_cfg5.py --mode strong --steps 30 2>&1 | grep '^{' | tee gpurun_out/cfg5_n2.jsonl
timeout 600 $R scripts/bench_cfg5.py --mode weak --steps 50 2>&1 | grep '^{' | tee -a gpurun_out/cfg5_n2.jsonl
timeout 600 python scripts/bench_cfg5.py --mode weak --steps 50 2>&1 | grep '^{' | tee -a gpurun_out/cfg5_n2.jsonl
timeout 600 $R scripts/bench_cfg5.py --mode weak --steps 50 --transport nccl 2>&1 | grep '^{' | tee -a gpurun_out/cfg5_n2.jsonl
