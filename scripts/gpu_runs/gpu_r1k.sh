for r in 1024x1024x128 1024x1024x256 1024x1024x64; do echo "== $r"; RES3=$r STEPS=30 timeout 300 python scripts/bench_variants.py 2>&1 | tail -6; done
