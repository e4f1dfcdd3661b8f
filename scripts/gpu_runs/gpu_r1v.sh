for v in 4 2; do echo "== LB_VEC=$v"; 
LB_VEC=$v timeout 200 python scripts/bench_configs.py "--only=stock lt.Obstacle D3Q27" 2>&1 | tail -1 | cut -c1-260
LB_VEC=$v timeout 200 python scripts/bench_configs.py "--only=TGV2D D2Q9 KBC" 2>&1 | tail -1 | cut -c1-260
LB_VEC=$v timeout 200 python scripts/bench_configs.py "--only=stock lt.Obstacle D2Q9" 2>&1 | tail -1 | cut -c1-260
done
for v in 1 2; do echo "== LB_VEC=$v"; LB_VEC=$v timeout 200 python scripts/bench_configs.py "--only=TGV3D D3Q27 kbc f64" 2>&1 | tail -1 | cut -c1-260; done
