set -x
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 600 python bench.py > gpurun_out/r1q_bench.json 2> gpurun_out/r1q_bench.err; tail -c 1500 gpurun_out/r1q_bench.json
timeout 900 python scripts/bench_configs.py > gpurun_out/r1q_configs.jsonl 2> gpurun_out/r1q_configs.err; tail -3 gpurun_out/r1q_configs.err
