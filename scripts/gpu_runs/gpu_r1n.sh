set -x
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -5
timeout 300 python scripts/bench_configs.py --quick --only=stock 2>&1 | tail -4
timeout 300 python scripts/bench_configs.py --quick --only=cfg3 2>&1 | tail -4
