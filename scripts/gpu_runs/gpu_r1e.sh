mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu.log
cat gpurun_out/pytest_gpu.log
timeout 900 python scripts/bench_configs.py --only=cfg4 --only=cfg1 > gpurun_out/configs_e.jsonl 2> gpurun_out/configs_e.err
cut -c1-220 gpurun_out/configs_e.jsonl; tail -3 gpurun_out/configs_e.err
