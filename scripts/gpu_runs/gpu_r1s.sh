set -x
timeout 1200 python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -5
timeout 300 python scripts/bench_configs.py --quick --only=rank3 2>&1 | tail -6
LB_VEC=2 timeout 300 python scripts/bench_configs.py --quick "--only=rank3 TGV3D D3Q27" 2>&1 | tail -2
