python scripts/obstacle_variants.py D3Q19 2>&1 | tail -7
python scripts/obstacle_variants.py D2Q9 2>&1 | tail -7
