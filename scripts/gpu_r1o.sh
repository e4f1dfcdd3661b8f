timeout 600 python -m pytest tests/test_gpu_slab_emulated.py -m gpu -x -q 2>&1 | tail -3
for r in 1024x1024x128; do echo "== $r"; RES3=$r STEPS=30 timeout 300 python scripts/bench_variants.py 2>&1 | tail -5; echo "-- no stage"; LB_HALO_NO_STAGE=1 RES3=$r STEPS=30 timeout 300 python scripts/bench_variants.py 2>&1 | tail -5; done
