mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -25
timeout 600 python scripts/bench_configs.py --only=cfg3 --only="stock" --only="cfg2 TGV3D D3Q19 bgk" | cut -c1-230
