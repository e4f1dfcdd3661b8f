set -x
timeout 1200 python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -5
timeout 300 python scripts/bench_configs.py --quick --only=rank3 2>&1 | tail -4
for v in 2 4; do LB_VEC=$v timeout 300 python scripts/bench_configs.py --quick "--only=rank3 TGV3D D3Q19" 2>&1 | tail -2; done
